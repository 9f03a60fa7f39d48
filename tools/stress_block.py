"""Repeats block-kernel renders and compares them byte for byte (races would show as flaky pixels)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import openglraytracer_b200 as rt
from test_gpu_parity import _far_scene
cam = rt.builtin_camera(0.0)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 40
with rt.Raytracer(0) as r:
    cases = [("far11", _far_scene(11.0), 320, 180), ("far3000", _far_scene(3000.0), 320, 180),
             ("scaled1k_d1", rt.scaled_scene(1024, max_depth=1), 640, 360),
             ("scaled16k", rt.scaled_scene(16384, max_depth=0), 960, 540)]
    for name, sc, w, h in cases:
        base = r.render(sc, cam, w, h, no_cull=4)
        for mode in (16, 48, 0):
            bad = 0
            for i in range(reps):
                img = r.render(sc, cam, w, h, no_cull=mode)
                n = int((img != base).any(axis=2).sum())
                if n:
                    bad += 1
                    ys, xs = np.nonzero((img != base).any(axis=2))
                    print("  ", name, "mode", mode, "rep", i, "diff pixels", n, "first", int(xs[0]), int(ys[0]))
            print(name, "mode", mode, "bad reps", bad, "of", reps)
