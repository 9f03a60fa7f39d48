"""Renders the far / tiny-object scenes of tests/test_gpu_parity.py under every no_cull switch, repeatedly, and reports any
pixel that differs from the per-ray bounding-test frame (all switches must give the same bytes)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import openglraytracer_b200 as rt
from test_gpu_parity import _far_scene
cam = rt.builtin_camera(0.0)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 30
with rt.Raytracer(0) as r:
    # warm the context the way the suite does: a big scene first
    warm = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
    if warm:
        r.render(rt.scaled_scene(warm, max_depth=0), cam, 256, 144)
    for half in (11.0, 3000.0):
        sc = _far_scene(half)
        ref = r.render(sc, cam, 320, 180, no_cull=4)
        for i in range(reps):
            for mode in (0, 1, 2, 4, 6, 8, 16, 16 | 32, 64, 8 | 64):
                img = r.render(sc, cam, 320, 180, no_cull=mode)
                d = (img != ref).any(axis=2)
                if d.any():
                    ys, xs = np.nonzero(d)
                    print("half", half, "rep", i, "mode", mode, "diff", int(d.sum()), "at", [(int(x), int(y)) for x, y in zip(xs[:6], ys[:6])],
                          ref[ys[0], xs[0]].tolist(), img[ys[0], xs[0]].tolist())
        print("half", half, "done")
