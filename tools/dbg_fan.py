"""Which pixels change when the fan planes of the block kernel are switched on (they must not)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import openglraytracer_b200 as rt
from test_gpu_parity import _far_scene
cam = rt.builtin_camera(0.0)
with rt.Raytracer(0) as r:
    for half in (11.0, 3000.0):
        sc = _far_scene(half)
        base = r.render(sc, cam, 320, 180)
        for mode in (16, 16 | 32, 8, 4):
            img = r.render(sc, cam, 320, 180, no_cull=mode)
            d = (img != base).any(axis=2)
            print("half", half, "mode", mode, "diff pixels", int(d.sum()))
            ys, xs = np.nonzero(d)
            for y, x in list(zip(ys, xs))[:10]:
                print("   ", x, y, base[y, x].tolist(), img[y, x].tolist())
