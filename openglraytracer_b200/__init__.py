"""openglraytracer_b200 — B200-native drop-in for OpenGLRaytracer's one hot path:
the per-pixel ray-scene kernel of raytrace_compute.glsl, as hand-written CUDA
for sm_100a behind the C ABI in include/rt_b200.h.  See DESIGN.md."""
from ._abi import (RT_MAX_DEPTH, RT_MAX_LIGHTS, Camera, Light, Material, Object, Options, RtError, Scene, Stats)
from .api import (PinnedImage, Raytracer, SceneData, builtin_camera, builtin_scene, fp32_peak_tflops,
                  kernel_launch_count, load_scene, orbit_camera, render, save_scene, scaled_scene, turntable_cameras, write_ppm,
                  write_raw)

__all__ = [
    "RT_MAX_DEPTH", "RT_MAX_LIGHTS", "Camera", "Light", "Material", "Object", "Options", "RtError", "Scene",
    "Stats", "PinnedImage", "Raytracer", "SceneData", "builtin_camera", "builtin_scene", "fp32_peak_tflops",
    "kernel_launch_count", "load_scene", "orbit_camera", "render", "save_scene", "scaled_scene", "turntable_cameras", "write_ppm",
    "write_raw",
]
