"""A/B of rt_options.no_cull switches on one workload: median kernel ms over N renders (CUDA events in the library,
rt_stats.kernel_ms of the plain launch).  python tools/ab_modes.py <n_objects|5> <w> <h> <depth> mode [mode ...]"""
import sys, os, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import openglraytracer_b200 as rt
n, w, h, depth = (int(x) for x in sys.argv[1:5])
modes = [int(x) for x in sys.argv[5:]] or [0]
sc = rt.scaled_scene(n, max_depth=depth) if n > 5 else rt.builtin_scene(0.0, depth)
cam = rt.builtin_camera(0.0)
r = rt.Raytracer()
r.upload_scene(sc)
ref = None
for m in modes:
    ms = []
    for i in range(12):
        img, st = r.render(None, cam, w, h, want_stats=True, no_cull=m)
        ms.append(st["kernel_ms"])
    if ref is None:
        ref = img
    print(json.dumps(dict(mode=m, kernel_ms=float(np.median(ms[2:])), min_ms=float(min(ms)), same=bool(np.array_equal(img, ref)),
                          box_tests=st["box_tests"], rays=st["rays"])))
