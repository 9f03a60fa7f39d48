"""Host-side mirror of the reference's dispatch interface for the ray-scene kernel.

The reference drives its compute shader with five GL calls per frame
(main.cpp:220-238): bind program, bind the RGBA8 output image, set the `time`
uniform, glDispatchCompute(W, H, 1), glFinish.  Here that is one call:

    rt = Raytracer()                       # GL context + texture  -> CUDA context + device buffers
    scene = builtin_scene(time)            # the shader's hard-coded objects/lights (glsl:199-321)
    cam = builtin_camera(time)             # the shader's orbit camera (glsl:334-364)
    rgba = rt.render(scene, cam, 1280, 720)   # uint8 [H, W, 4], row 0 = bottom scanline, A = 0

Everything goes through the C ABI (include/rt_b200.h); numpy only holds the
host pixels.  No CPU fallback exists.
"""
import ctypes as C
import os

import numpy as np

from . import _abi
from ._abi import Camera, Light, Material, Object, Options, Scene, Stats, check


class SceneData:
    """Owns the ctypes arrays an rt_scene points at (objects[], lights[], max depth)."""

    def __init__(self, objects, lights, max_depth=0):
        self.objects = objects
        self.lights = lights
        self.scene = Scene(C.cast(objects, C.POINTER(Object)), len(objects),
                           C.cast(lights, C.POINTER(Light)), len(lights), int(max_depth))

    @property
    def max_depth(self):
        return self.scene.max_depth

    @max_depth.setter
    def max_depth(self, d):
        self.scene.max_depth = int(d)

    def with_depth(self, d):
        self.max_depth = d
        return self

    @property
    def objects_count(self):
        return self.scene.objects_count

    def pinned(self):
        """The same scene with its objects[] in page-locked host memory (rt_host_alloc): large scenes then go up
        over PCIe straight from the array at full rate (the device prologue copies them in one DMA, no staging)."""
        n = max(1, len(self.objects))
        ptr = C.c_void_p()
        check(_abi.lib().rt_host_alloc(C.byref(ptr), n * C.sizeof(Object)), "rt_host_alloc")
        objs = (Object * n).from_address(ptr.value)
        C.memmove(objs, self.objects, len(self.objects) * C.sizeof(Object))
        out = SceneData(objs, self.lights, self.max_depth)
        out.scene.objects_count = self.scene.objects_count
        out.scene.lights_count = self.scene.lights_count
        out._pinned_ptr = ptr
        return out

    def free_pinned(self):
        ptr = getattr(self, "_pinned_ptr", None)
        if ptr:
            self.objects = None
            _abi.lib().rt_host_free(ptr)
            self._pinned_ptr = None


def builtin_scene(time=0.0, max_depth=0):
    """objects[5], lights[3] of the shader at uniform `time` (glsl:199-224, 261-321).
    max_depth defaults to the shipped MAX_RAYTRACE_DEPTH = 0 (glsl:22)."""
    objs = (Object * 5)()
    lights = (Light * 3)()
    sc = Scene()
    check(_abi.lib().rt_builtin_scene(C.c_float(time), objs, lights, C.byref(sc)), "rt_builtin_scene")
    return SceneData(objs, lights, max_depth)


def builtin_camera(time=0.0):
    """The shader's orbiting camera at uniform `time` (glsl:334-364)."""
    cam = Camera()
    check(_abi.lib().rt_builtin_camera(C.c_float(time), C.byref(cam)), "rt_builtin_camera")
    return cam


def orbit_camera(speed):
    """Camera from the shader's `speed` variable (glsl:343-358); turntable poses."""
    cam = Camera()
    check(_abi.lib().rt_orbit_camera(C.c_float(speed), C.byref(cam)), "rt_orbit_camera")
    return cam


def turntable_cameras(n):
    """BASELINE config 5: n poses, speed_k = 0.5 + 2*pi*k/n."""
    cams = (Camera * n)()
    for k in range(n):
        check(_abi.lib().rt_orbit_camera(C.c_float(np.float32(0.5 + 2.0 * np.pi * k / n)), C.byref(cams[k])),
              "rt_orbit_camera")
    return cams


def scaled_scene(n_objects, seed=0x5CE7E001, max_depth=0):
    """BASELINE config 4: the shader's primitive types replicated to n_objects."""
    objs = (Object * n_objects)()
    lights = (Light * 3)()
    sc = Scene()
    check(_abi.lib().rt_scaled_scene(n_objects, C.c_uint64(seed), objs, lights, C.byref(sc)), "rt_scaled_scene")
    return SceneData(objs, lights, max_depth)


def write_ppm(path, rgba):
    """Replaces the fullscreen-quad blit (main.cpp:240-260): upright binary PPM."""
    rgba = np.ascontiguousarray(rgba, dtype=np.uint8)
    h, w = rgba.shape[:2]
    check(_abi.lib().rt_write_ppm(path.encode(), rgba.ctypes.data, w, h), "rt_write_ppm")


def write_raw(path, rgba):
    rgba = np.ascontiguousarray(rgba, dtype=np.uint8)
    h, w = rgba.shape[:2]
    check(_abi.lib().rt_write_raw(path.encode(), rgba.ctypes.data, w, h), "rt_write_raw")


class PinnedImage:
    """Page-locked host framebuffer (rt_host_alloc) viewed as a numpy array."""

    def __init__(self, shape):
        self.shape = tuple(shape)
        self.nbytes = int(np.prod(self.shape))
        self.ptr = C.c_void_p()
        check(_abi.lib().rt_host_alloc(C.byref(self.ptr), self.nbytes), "rt_host_alloc")
        buf = (C.c_uint8 * self.nbytes).from_address(self.ptr.value)
        self.array = np.frombuffer(buf, dtype=np.uint8).reshape(self.shape)

    def free(self):
        if self.ptr:
            self.array = None
            _abi.lib().rt_host_free(self.ptr)
            self.ptr = C.c_void_p()


def render(scene, camera_uniforms, width, height):
    """The drop-in call: rt_render(scene, camera_uniforms, out_rgba, width, height)."""
    out = np.empty((height, width, 4), np.uint8)
    sc = scene.scene if isinstance(scene, SceneData) else scene
    check(_abi.lib().rt_render(C.byref(sc), C.byref(camera_uniforms), out.ctypes.data, width, height), "rt_render")
    return out


class Raytracer:
    """An rt_context: device buffers + stream on one GPU (replaces the GL context
    and output texture, main.cpp:47-76, 152-159)."""

    def __init__(self, device=-1):
        self._lib = _abi.lib()
        self._ctx = C.c_void_p()
        check(self._lib.rt_create(C.byref(self._ctx), device), "rt_create")

    def close(self):
        if self._ctx:
            self._lib.rt_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def set_prologue(self, mode):
        """Where the per-object prologue runs: "auto" (device from 2048 objects), "host", "device"."""
        code = {"auto": 0, "host": 1, "device": 2}[mode]
        check(self._lib.rt_set_prologue(self._ctx, code), "rt_set_prologue")

    def synchronize(self):
        check(self._lib.rt_synchronize(self._ctx), "rt_synchronize")

    @staticmethod
    def _options(spp, seed, y0, y1, strict, out_is_device, band_only, async_, stream, stats, no_cull=False, host_store=0,
                 done_flag=None, done_value=0):
        o = Options()
        _abi.lib().rt_options_default(C.byref(o))
        o.spp, o.seed, o.y0, o.y1 = int(spp), int(seed), int(y0), int(y1)
        o.strict, o.out_is_device, o.out_band_only = int(strict), int(out_is_device), int(band_only)
        o.async_ = int(async_)
        o.stream = stream
        o.no_cull = int(no_cull)
        o.host_store = int(host_store)
        if done_flag:
            o.done_flag = C.c_void_p(int(done_flag))
            o.done_value = int(done_value) & 0xFFFFFFFF
        if stats is not None:
            o.stats = C.pointer(stats)
        return o

    def upload_scene(self, scene):
        """Host prologue + upload (per-object matrices hoisted out of glsl:650-652, 718)."""
        sc = scene.scene if isinstance(scene, SceneData) else scene
        check(self._lib.rt_upload_scene(self._ctx, C.byref(sc)), "rt_upload_scene")

    def render(self, scene, camera_uniforms, width, height, *, spp=1, seed=0x5EED0001, y0=0, y1=0,
               strict=False, out=None, out_device_ptr=None, band_only=False, async_=False, stream=None,
               want_stats=False, no_cull=False, host_store=0, done_flag=None, done_value=0):
        """Render one frame (camera_uniforms: Camera) or a batch (ctypes array of Camera).

        scene=None re-uses the last uploaded scene.  Host output: numpy uint8
        [H, W, 4] (or [n, H, W, 4]); with band_only the H axis is y1-y0.  Device
        output: pass out_device_ptr (e.g. torch_tensor.data_ptr()), returns None.
        done_flag / done_value (device output only): rt_options.done_flag, the completion word the launch's
        last CTA writes (multi-GPU frame mode, see multigpu.PeerFrame).
        """
        batch = isinstance(camera_uniforms, C.Array)
        n = len(camera_uniforms) if batch else 1
        cams = camera_uniforms if batch else C.byref(camera_uniforms)
        cams = C.cast(cams, C.POINTER(Camera))
        yy1 = y1 if (0 < y1 <= height) else height
        rows = (yy1 - max(y0, 0)) if band_only else height
        stats = Stats() if want_stats else None
        if out_device_ptr is not None:
            opts = self._options(spp, seed, y0, y1, strict, 1, band_only, async_, stream, stats, no_cull,
                                 done_flag=done_flag, done_value=done_value)
            target = C.c_void_p(int(out_device_ptr))
            result = None
        else:
            if out is None:
                out = np.empty((n, rows, width, 4) if batch else (rows, width, 4), np.uint8)
            if out.nbytes != n * rows * width * 4 or not out.flags["C_CONTIGUOUS"] or out.dtype != np.uint8:
                raise ValueError("out must be a C-contiguous uint8 array of %d bytes" % (n * rows * width * 4))
            opts = self._options(spp, seed, y0, y1, strict, 0, band_only, False, stream, stats, no_cull, host_store)
            target = C.c_void_p(out.ctypes.data)
            result = out
        if scene is not None:
            self.upload_scene(scene)
        check(self._lib.rt_render_uploaded(self._ctx, cams, n, target, width, height, C.byref(opts)),
              "rt_render_uploaded")
        if want_stats:
            return result, stats.as_dict()
        return result


def render_block_noalloc(raytracer, camera_uniforms, width, height, out_device_ptr, *, spp=1, seed=0x5EED0001, y0=0, y1=0,
                         band_only=False, async_=False, stream=None, no_cull=0):
    """rt_render_block_noalloc: the uploaded scene through the block kernel with all scratch in the library's
    file-scope device array -- the call allocates nothing (no counterpart in the reference; see include/rt_b200.h)."""
    batch = isinstance(camera_uniforms, C.Array)
    n = len(camera_uniforms) if batch else 1
    cams = C.cast(camera_uniforms if batch else C.byref(camera_uniforms), C.POINTER(Camera))
    opts = Raytracer._options(spp, seed, y0, y1, False, 1, band_only, async_, stream, None, no_cull)
    check(_abi.lib().rt_render_block_noalloc(raytracer._ctx, cams, n, C.c_void_p(int(out_device_ptr)), width, height,
                                             C.byref(opts)), "rt_render_block_noalloc")


def fp32_peak_tflops(device=-1, iters=4096):
    """FP32 FMA-chain microbenchmark (2 flop/FMA): roofline denominator check."""
    tf = C.c_double()
    ms = C.c_float()
    check(_abi.lib().rt_fp32_peak_tflops(device, iters, C.byref(tf), C.byref(ms)), "rt_fp32_peak_tflops")
    return tf.value, ms.value


def wait_flags(flags_device_ptr, n, value, stream=None, timeout_ms=2000):
    """One warp on `stream` waits until the n 32-bit words at flags_device_ptr have reached `value` (rt_wait_flags)."""
    check(_abi.lib().rt_wait_flags(C.c_void_p(int(flags_device_ptr)), int(n), int(value) & 0xFFFFFFFF, int(timeout_ms),
                                   C.c_void_p(stream) if stream else None), "rt_wait_flags")


def wait_flags_timed_out():
    return _abi.lib().rt_wait_flags_status() == 1


def kernel_launch_count():
    return int(_abi.lib().rt_kernel_launch_count())


def save_scene(path, scene, camera=None):
    """Write the scene (and optionally the camera uniforms) as an `rtscene 1` text file:
    the shader's hard-coded globals (glsl:22, 199-224, 261-321, 334-364) on the wire."""
    sc = scene.scene if isinstance(scene, SceneData) else scene
    cam = C.byref(camera) if camera is not None else None
    check(_abi.lib().rt_scene_save(os.fsencode(path), C.byref(sc), cam), "rt_scene_save")


def load_scene(path):
    """Read an `rtscene 1` file: returns (SceneData, Camera or None).  The arrays are copied
    out of the loader's storage, so the result owns its memory."""
    L = _abi.lib()
    h = C.c_void_p()
    check(L.rt_scene_load(os.fsencode(path), C.byref(h)), "rt_scene_load")
    try:
        sc = L.rt_scene_file_scene(h).contents
        objs = (Object * max(1, sc.objects_count))()
        lights = (Light * max(1, sc.lights_count))()
        for i in range(sc.objects_count):
            C.memmove(C.byref(objs[i]), C.byref(sc.objects[i]), C.sizeof(Object))
        for j in range(sc.lights_count):
            C.memmove(C.byref(lights[j]), C.byref(sc.lights[j]), C.sizeof(Light))
        data = SceneData(objs, lights, sc.max_depth)
        data.scene.objects_count = sc.objects_count
        data.scene.lights_count = sc.lights_count
        cp = L.rt_scene_file_camera(h)
        cam = None
        if cp:
            cam = Camera()
            C.memmove(C.byref(cam), cp, C.sizeof(Camera))
        return data, cam
    finally:
        L.rt_scene_file_free(h)
