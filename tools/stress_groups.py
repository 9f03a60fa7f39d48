"""Random scenes and cameras through the object groups (rt_groups.cu) and the cluster cones: the shipped frame against
the per-ray loops (no_cull = 4: no packet tests at all) and against the groups switched off, byte for byte."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import openglraytracer_b200 as rt

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 6
bad = 0
with rt.Raytracer(0) as r:
    for k in range(reps):
        for n, depth, w, h in ((2100, 0, 320, 180), (5000, 1, 240, 136), (20000, 0, 320, 180), (9000, 2, 160, 90),
                               (70000, 0, 480, 270), (300, 0, 320, 180), (3000, 0, 640, 360)):
            sc = rt.scaled_scene(n, seed=0xC0FFEE00 + 977 * k + n, max_depth=depth)
            cam = rt.builtin_camera(0.37 * k + 0.11 * depth)
            a = r.render(sc, cam, w, h)
            for mode in (4, 256, 512, 8, 16):
                b = r.render(sc, cam, w, h, no_cull=mode)
                d = int((a != b).any(axis=2).sum())
                if d:
                    bad += 1
                    print("DIFF", k, n, depth, "mode", mode, "pixels", d)
print("cases", reps * 7 * 5, "bad", bad)
