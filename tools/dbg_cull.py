"""Debug helper: where does the bounding-sphere path differ from the plain one?"""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import openglraytracer_b200 as rt

r = rt.Raytracer()
cam = rt.builtin_camera(0.0)
for n, w, h, depth in [(200, 320, 180, 0), (200, 320, 180, 1), (200, 320, 180, 2), (1024, 480, 270, 1), (4096, 480, 270, 0)]:
    sc = rt.scaled_scene(n, max_depth=depth)
    on, s1 = r.render(sc, cam, w, h, want_stats=True)
    off, s0 = r.render(sc, cam, w, h, want_stats=True, no_cull=True)
    on2 = r.render(sc, cam, w, h)
    off2 = r.render(sc, cam, w, h, no_cull=True)
    d = np.abs(on.astype(int) - off.astype(int)).max(axis=2)
    ys, xs = np.nonzero(d)
    print(n, depth, "diff px", len(ys), "max", d.max(), "plain-vs-count on", (on != on2).sum(), "off", (off != off2).sum(),
          "rays", s1["rays"], s0["rays"], "hit", s1["hit_nodes"], s0["hit_nodes"])
    for y, x in list(zip(ys, xs))[:8]:
        print("   ", x, y, on[y, x], off[y, x])
