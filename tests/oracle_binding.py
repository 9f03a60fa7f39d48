"""ctypes binding of the CPU oracle (oracle/librt_oracle.so).

TEST INFRASTRUCTURE: imported only by tests/, bench.py's cpu_baseline / --impl
reference legs and __graft_entry__.smoke().  The product package never imports it.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "librt_oracle.so")


class Material(C.Structure):
    _fields_ = [("ambient", C.c_float * 4), ("diffuse", C.c_float * 4), ("specular", C.c_float * 4),
                ("shininess", C.c_float), ("emissive", C.c_float * 4), ("reflectivity", C.c_float),
                ("transparency", C.c_float), ("refraction_index", C.c_float)]


class Object(C.Structure):
    _fields_ = [("box_mins", C.c_float * 3), ("box_maxs", C.c_float * 3), ("sphere_radius", C.c_float),
                ("position", C.c_float * 3), ("angles", C.c_float * 3), ("material", Material)]


class Light(C.Structure):
    _fields_ = [("position", C.c_float * 3), ("ambient", C.c_float * 4), ("diffuse", C.c_float * 4),
                ("specular", C.c_float * 4)]


class Scene(C.Structure):
    _fields_ = [("objects", C.POINTER(Object)), ("objects_count", C.c_int32),
                ("lights", C.POINTER(Light)), ("lights_count", C.c_int32), ("max_depth", C.c_int32)]


class Camera(C.Structure):
    _fields_ = [("position", C.c_float * 3), ("angles", C.c_float * 3), ("v_fov", C.c_float),
                ("aspect", C.c_float), ("near_plane", C.c_float), ("far_plane", C.c_float)]


class OracleStats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("rays", "nodes", "hit_nodes", "shadow_rays", "box_tests",
                                          "sphere_tests", "red_pixels", "loop_iterations")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


class OracleOptions(C.Structure):
    _fields_ = [("spp", C.c_int32), ("seed", C.c_uint32), ("y0", C.c_int32), ("y1", C.c_int32),
                ("as_written", C.c_int32), ("threads", C.c_int32), ("pixel_list", C.POINTER(C.c_int32)),
                ("n_pixels", C.c_int32), ("out_float", C.POINTER(C.c_float)), ("stats", C.POINTER(OracleStats)),
                ("out_nodes", C.POINTER(C.c_int32))]


_lib = None


def build_oracle(force=False):
    if force or not os.path.exists(ORACLE_SO) or any(
            os.path.getmtime(os.path.join(ORACLE_DIR, f)) > os.path.getmtime(ORACLE_SO)
            for f in ("rt_oracle.cpp", "rt_oracle.h", "glsl_shim.h")):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "-s"])
    return ORACLE_SO


def lib():
    global _lib
    if _lib is None:
        build_oracle()
        _lib = C.CDLL(ORACLE_SO)
        _lib.oracle_render.restype = C.c_int
        _lib.oracle_render.argtypes = [C.POINTER(Scene), C.POINTER(Camera), C.c_void_p, C.c_int, C.c_int,
                                       C.POINTER(OracleOptions)]
        _lib.oracle_builtin_scene.argtypes = [C.c_float, C.POINTER(Object), C.POINTER(Light), C.POINTER(Scene)]
        _lib.oracle_builtin_camera.argtypes = [C.c_float, C.POINTER(Camera)]
        _lib.oracle_orbit_camera.argtypes = [C.c_float, C.POINTER(Camera)]
        _lib.oracle_unorm8.restype = C.c_uint8
        _lib.oracle_unorm8.argtypes = [C.c_float]
        _lib.oracle_num_threads.restype = C.c_int
        _lib.oracle_refract.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_void_p]
    return _lib


class SceneHolder:
    """Keeps the ctypes arrays alive next to the Scene that points at them."""

    def __init__(self, objects, lights, max_depth=0):
        self.objects = objects
        self.lights = lights
        self.scene = Scene(C.cast(objects, C.POINTER(Object)), len(objects),
                           C.cast(lights, C.POINTER(Light)), len(lights), max_depth)

    def with_depth(self, depth):
        self.scene.max_depth = depth
        return self


def builtin_scene(time=0.0, max_depth=0):
    objs = (Object * 5)()
    lights = (Light * 3)()
    sc = Scene()
    lib().oracle_builtin_scene(C.c_float(time), objs, lights, C.byref(sc))
    return SceneHolder(objs, lights, max_depth)


def builtin_camera(time=0.0):
    cam = Camera()
    lib().oracle_builtin_camera(C.c_float(time), C.byref(cam))
    return cam


def orbit_camera(speed):
    cam = Camera()
    lib().oracle_orbit_camera(C.c_float(speed), C.byref(cam))
    return cam


def render(scene, camera, width, height, spp=1, seed=0x5EED0001, y0=0, y1=0, as_written=False,
           threads=0, pixels=None, want_float=False, want_stats=False, want_nodes=False):
    """scene: SceneHolder-like (has .scene) or a Scene with the oracle's layout."""
    sc = scene.scene if hasattr(scene, "scene") else scene
    # accept the product binding's structs too (identical layout)
    sc_p = C.cast(C.pointer(sc), C.POINTER(Scene))
    cam_p = C.cast(C.pointer(camera), C.POINTER(Camera))
    opt = OracleOptions()
    opt.spp, opt.seed, opt.y0, opt.y1 = spp, seed, y0, y1
    opt.as_written, opt.threads = int(as_written), threads
    if pixels is not None:
        pix = np.ascontiguousarray(pixels, dtype=np.int32).reshape(-1, 2)
        opt.pixel_list = pix.ctypes.data_as(C.POINTER(C.c_int32))
        opt.n_pixels = pix.shape[0]
        out = np.zeros((pix.shape[0], 4), np.uint8)
        fl = np.zeros((pix.shape[0], 3), np.float32)
    else:
        out = np.zeros((height, width, 4), np.uint8)
        fl = np.zeros((height, width, 3), np.float32)
    if want_float:
        opt.out_float = fl.ctypes.data_as(C.POINTER(C.c_float))
    nodes = np.zeros(out.shape[:-1], np.int32)
    if want_nodes:
        opt.out_nodes = nodes.ctypes.data_as(C.POINTER(C.c_int32))
    st = OracleStats()
    if want_stats:
        opt.stats = C.pointer(st)
    rc = lib().oracle_render(sc_p, cam_p, out.ctypes.data, width, height, C.byref(opt))
    if rc != 0:
        raise RuntimeError("oracle_render failed: %d" % rc)
    res = [out]
    if want_float:
        res.append(fl)
    if want_stats:
        res.append(st.as_dict())
    if want_nodes:
        res.append(nodes)
    return res[0] if len(res) == 1 else tuple(res)


def probe_pixel(scene, camera, width, height, px, py):
    sc = scene.scene if hasattr(scene, "scene") else scene
    d = (C.c_float * 3)()
    p = (C.c_float * 3)()
    n = (C.c_float * 3)()
    obj = C.c_int()
    t = C.c_float()
    inside = C.c_int()
    lib().oracle_probe_pixel(C.byref(sc), C.byref(camera), width, height, px, py, d, C.byref(obj), C.byref(t), p, n,
                             C.byref(inside))
    return dict(dir=np.array(d[:]), obj=obj.value, t=t.value, p=np.array(p[:]), n=np.array(n[:]),
                inside=bool(inside.value))


def probe_ray(scene, start, direction):
    sc = scene.scene if hasattr(scene, "scene") else scene
    s = (C.c_float * 3)(*start)
    d = (C.c_float * 3)(*direction)
    p = (C.c_float * 3)()
    n = (C.c_float * 3)()
    obj = C.c_int()
    t = C.c_float()
    inside = C.c_int()
    lib().oracle_probe_ray(C.byref(sc), s, d, C.byref(obj), C.byref(t), p, n, C.byref(inside))
    return dict(obj=obj.value, t=t.value, p=np.array(p[:]), n=np.array(n[:]), inside=bool(inside.value))
