"""First-contact GPU check: renders several configs with the strict and fast
kernels, compares with the oracle, prints diff histograms and timings."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import openglraytracer_b200 as rt  # noqa: E402
import oracle_binding as ob  # noqa: E402
from parity import compare  # noqa: E402


def main():
    out = {}
    r = rt.Raytracer()
    tf, ms = rt.fp32_peak_tflops(iters=8192)
    print("fp32 peak TFLOP/s", tf, "ms", ms)
    out["fp32_peak_tflops"] = tf
    for (t, depth, w, h) in [(0.0, 0, 1280, 720), (0.0, 2, 1280, 720), (2.5, 4, 1280, 720), (3.7, 4, 640, 360),
                             (0.0, 8, 320, 180)]:
        sc = rt.builtin_scene(t, depth)
        cam = rt.builtin_camera(t)
        ref, ost = ob.render(sc, cam, w, h, want_stats=True)
        for strict in (True, False):
            img, st = r.render(sc, cam, w, h, strict=strict, want_stats=True)
            rep = compare(img, ref)
            key = "builtin t=%g d=%d %dx%d %s" % (t, depth, w, h, "strict" if strict else "fast")
            rep.update(kernel_ms=st["kernel_ms"], rays=st["rays"], oracle_rays=ost["rays"],
                       box_tests=st["box_tests"], oracle_box_tests=ost["box_tests"], red=st["red_pixels"],
                       oracle_red=ost["red_pixels"])
            print(key, json.dumps(rep))
            out[key] = rep
    # scaled scenes, small frame full parity
    for n in (64, 1024, 4096):
        sc = rt.scaled_scene(n, max_depth=1)
        cam = rt.builtin_camera(0.0)
        w, h = 256, 144
        ref, ost = ob.render(sc, cam, w, h, want_stats=True)
        for strict in (True, False):
            img, st = r.render(sc, cam, w, h, strict=strict, want_stats=True)
            rep = compare(img, ref)
            key = "scaled n=%d d=1 %dx%d %s" % (n, w, h, "strict" if strict else "fast")
            rep.update(kernel_ms=st["kernel_ms"], rays=st["rays"], oracle_rays=ost["rays"],
                       box_tests=st["box_tests"], oracle_box_tests=ost["box_tests"])
            print(key, json.dumps(rep))
            out[key] = rep
    # perf probes
    for n in (1024, 4096, 16384, 65536):
        sc = rt.scaled_scene(n, max_depth=0)
        cam = rt.builtin_camera(0.0)
        w, h = (1920, 1080) if n <= 4096 else (480, 270)
        r.render(sc, cam, w, h)
        img, st = r.render(sc, cam, w, h, want_stats=True)
        flops = 59.0 * st["box_tests"] + 41.0 * st["sphere_tests"]
        key = "perf scaled n=%d d=0 %dx%d" % (n, w, h)
        rep = dict(kernel_ms=st["kernel_ms"], rays=st["rays"], box_tests=st["box_tests"],
                   sphere_tests=st["sphere_tests"], tflops=flops / (st["kernel_ms"] * 1e-3) / 1e12,
                   mrays=st["rays"] / (st["kernel_ms"] * 1e-3) / 1e6)
        print(key, json.dumps(rep))
        out[key] = rep
    for depth in (0, 4):
        sc = rt.builtin_scene(0.0, depth)
        cam = rt.builtin_camera(0.0)
        r.render(sc, cam, 1920, 1080)
        img, st = r.render(sc, cam, 1920, 1080, want_stats=True)
        key = "perf builtin d=%d 1080p" % depth
        rep = dict(kernel_ms=st["kernel_ms"], total_ms=st["total_ms"], rays=st["rays"],
                   mrays=st["rays"] / (st["kernel_ms"] * 1e-3) / 1e6)
        print(key, json.dumps(rep))
        out[key] = rep
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "gpu_check.json"), "w") as f:
        json.dump(out, f, indent=1)
    rt.write_ppm(os.path.join(ROOT, "gpurun_out", "builtin_1080p_d4.ppm"), img)


if __name__ == "__main__":
    main()
