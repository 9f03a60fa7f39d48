"""ctypes view of the C ABI in include/rt_b200.h (librt_b200.so).

The product has NO CPU fallback: if the shared library is missing this module
raises at import, and every render call fails loudly when no CUDA device is
usable (RT_ERR_NO_DEVICE).  Nothing here touches oracle/.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RT_B200_LIB", os.path.join(_HERE, "librt_b200.so"))   # override: kernel experiments only

RT_OK = 0
RT_ERR_BAD_ARG = -1
RT_ERR_CUDA = -2
RT_ERR_NO_DEVICE = -3
RT_ERR_UNSUPPORTED = -5
RT_NOALLOC_MAX_OBJECTS = 2048
RT_NOALLOC_MAX_FRAMES = 16
RT_MAX_DEPTH = 15
RT_MAX_LIGHTS = 64


class Material(C.Structure):
    """glsl:56-69 struct Material."""
    _fields_ = [("ambient", C.c_float * 4), ("diffuse", C.c_float * 4), ("specular", C.c_float * 4),
                ("shininess", C.c_float), ("emissive", C.c_float * 4), ("reflectivity", C.c_float),
                ("transparency", C.c_float), ("refraction_index", C.c_float)]


class Object(C.Structure):
    """glsl:244-258 struct Object (Box and Sphere inlined)."""
    _fields_ = [("box_mins", C.c_float * 3), ("box_maxs", C.c_float * 3), ("sphere_radius", C.c_float),
                ("position", C.c_float * 3), ("angles", C.c_float * 3), ("material", Material)]


class Light(C.Structure):
    """glsl:190-196 struct Light."""
    _fields_ = [("position", C.c_float * 3), ("ambient", C.c_float * 4), ("diffuse", C.c_float * 4),
                ("specular", C.c_float * 4)]


class Scene(C.Structure):
    _fields_ = [("objects", C.POINTER(Object)), ("objects_count", C.c_int32),
                ("lights", C.POINTER(Light)), ("lights_count", C.c_int32), ("max_depth", C.c_int32)]


class Camera(C.Structure):
    """glsl:36-50 struct Camera."""
    _fields_ = [("position", C.c_float * 3), ("angles", C.c_float * 3), ("v_fov", C.c_float),
                ("aspect", C.c_float), ("near_plane", C.c_float), ("far_plane", C.c_float)]


class Stats(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("nodes", C.c_uint64), ("hit_nodes", C.c_uint64),
                ("shadow_rays", C.c_uint64), ("box_tests", C.c_uint64), ("sphere_tests", C.c_uint64),
                ("red_pixels", C.c_uint64), ("rays_not_traversed", C.c_uint64), ("kernel_ms", C.c_float), ("total_ms", C.c_float),
                ("cull_tests", C.c_uint64), ("warp_iters", C.c_uint64),
                ("packet_tests", C.c_uint64), ("group_passes", C.c_uint64), ("loop_passes", C.c_uint64),
                ("candidates", C.c_uint64), ("shadow_group_passes", C.c_uint64), ("shaded_lights", C.c_uint64)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class Options(C.Structure):
    _fields_ = [("spp", C.c_int32), ("seed", C.c_uint32), ("y0", C.c_int32), ("y1", C.c_int32),
                ("out_is_device", C.c_int32), ("out_band_only", C.c_int32), ("device", C.c_int32),
                ("strict", C.c_int32), ("async_", C.c_int32), ("stream", C.c_void_p),
                ("stats", C.POINTER(Stats)), ("host_store", C.c_int32), ("no_cull", C.c_int32),
                ("done_flag", C.c_void_p), ("done_value", C.c_uint32)]


# Every symbol include/rt_b200.h declares (tests check the library exports all of them).
EXPORTS = [
    "rt_abi_version", "rt_last_error_string", "rt_kernel_launch_count", "rt_create", "rt_destroy",
    "rt_synchronize", "rt_options_default", "rt_host_alloc", "rt_host_free", "rt_render", "rt_render_ex",
    "rt_render_batch", "rt_upload_scene", "rt_render_uploaded", "rt_render_block_noalloc", "rt_builtin_scene", "rt_builtin_camera",
    "rt_orbit_camera", "rt_scaled_scene", "rt_write_ppm", "rt_write_raw", "rt_fp32_peak_tflops",
    "rt_ipc_export", "rt_ipc_import", "rt_ipc_close", "rt_device_alloc", "rt_device_free",
    "rt_wait_flags", "rt_wait_flags_status", "rt_write_flag", "rt_set_prologue", "rt_scene_save", "rt_scene_load", "rt_scene_file_scene", "rt_scene_file_camera", "rt_scene_file_free",
]

_lib = None


def lib():
    """Load librt_b200.so (built in-tree by __graft_entry__.build() / csrc/Makefile)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "openglraytracer_b200: %s is missing. Build it with `make -C openglraytracer_b200/csrc` "
            "(or __graft_entry__.build()). There is no CPU fallback." % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp = C.c_void_p
    L.rt_abi_version.restype = C.c_int
    L.rt_last_error_string.restype = C.c_char_p
    L.rt_kernel_launch_count.restype = C.c_uint64
    L.rt_create.argtypes = [C.POINTER(vp), C.c_int]
    L.rt_destroy.argtypes = [vp]
    L.rt_synchronize.argtypes = [vp]
    L.rt_options_default.argtypes = [C.POINTER(Options)]
    L.rt_options_default.restype = None
    L.rt_host_alloc.argtypes = [C.POINTER(vp), C.c_size_t]
    L.rt_host_free.argtypes = [vp]
    L.rt_render.argtypes = [C.POINTER(Scene), C.POINTER(Camera), vp, C.c_int, C.c_int]
    L.rt_render_ex.argtypes = [vp, C.POINTER(Scene), C.POINTER(Camera), vp, C.c_int, C.c_int, C.POINTER(Options)]
    L.rt_render_batch.argtypes = [vp, C.POINTER(Scene), C.POINTER(Camera), C.c_int, vp, C.c_int, C.c_int,
                                  C.POINTER(Options)]
    L.rt_upload_scene.argtypes = [vp, C.POINTER(Scene)]
    L.rt_render_uploaded.argtypes = [vp, C.POINTER(Camera), C.c_int, vp, C.c_int, C.c_int, C.POINTER(Options)]
    L.rt_render_block_noalloc.argtypes = [vp, C.POINTER(Camera), C.c_int, vp, C.c_int, C.c_int, C.POINTER(Options)]
    L.rt_builtin_scene.argtypes = [C.c_float, C.POINTER(Object), C.POINTER(Light), C.POINTER(Scene)]
    L.rt_builtin_camera.argtypes = [C.c_float, C.POINTER(Camera)]
    L.rt_orbit_camera.argtypes = [C.c_float, C.POINTER(Camera)]
    L.rt_scaled_scene.argtypes = [C.c_int, C.c_uint64, C.POINTER(Object), C.POINTER(Light), C.POINTER(Scene)]
    L.rt_write_ppm.argtypes = [C.c_char_p, vp, C.c_int, C.c_int]
    L.rt_write_raw.argtypes = [C.c_char_p, vp, C.c_int, C.c_int]
    L.rt_fp32_peak_tflops.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_float)]
    L.rt_ipc_export.argtypes = [vp, C.c_char_p]
    L.rt_ipc_import.argtypes = [C.c_char_p, C.POINTER(vp)]
    L.rt_ipc_close.argtypes = [vp]
    L.rt_device_alloc.argtypes = [C.POINTER(vp), C.c_size_t]
    L.rt_device_free.argtypes = [vp]
    L.rt_set_prologue.argtypes = [vp, C.c_int]
    L.rt_wait_flags.argtypes = [vp, C.c_int, C.c_uint32, C.c_int, vp]
    L.rt_wait_flags_status.argtypes = []
    L.rt_write_flag.argtypes = [vp, C.c_uint32, vp]
    L.rt_scene_save.argtypes = [C.c_char_p, C.POINTER(Scene), C.POINTER(Camera)]
    L.rt_scene_load.argtypes = [C.c_char_p, C.POINTER(vp)]
    L.rt_scene_file_scene.argtypes = [vp]
    L.rt_scene_file_camera.argtypes = [vp]
    L.rt_scene_file_free.argtypes = [vp]
    for name in EXPORTS:
        fn = getattr(L, name)
        if name not in ("rt_last_error_string", "rt_kernel_launch_count", "rt_options_default",
                        "rt_scene_file_scene", "rt_scene_file_camera", "rt_scene_file_free"):
            fn.restype = C.c_int
    L.rt_scene_file_scene.restype = C.POINTER(Scene)
    L.rt_scene_file_camera.restype = C.POINTER(Camera)
    L.rt_scene_file_free.restype = None
    _lib = L
    return L


class RtError(RuntimeError):
    def __init__(self, code, where):
        msg = lib().rt_last_error_string()
        super().__init__("%s failed (%d): %s" % (where, code, msg.decode() if msg else ""))
        self.code = code


def check(code, where):
    if code != RT_OK:
        raise RtError(code, where)
