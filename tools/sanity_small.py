"""Small renders through every kernel variant (compute-sanitizer target): small resident, large resident
(bounding tests + light records), streaming, no-cull, strict, device prologue, spp, batch, bands."""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import openglraytracer_b200 as rt
r = rt.Raytracer()
cam = rt.builtin_camera(0.0)
tot = 0
for name, sc, w, h, kw in [
        ("builtin d3", rt.builtin_scene(0.3, 3), 96, 54, {}),
        ("builtin strict", rt.builtin_scene(0.3, 2), 64, 36, {"strict": True}),
        ("builtin spp4 band", rt.builtin_scene(0.3, 1), 64, 36, {"spp": 4, "y0": 5, "y1": 30, "band_only": True}),
        ("scaled 200 d1", rt.scaled_scene(200, max_depth=1), 72, 40, {}),
        ("scaled 200 nocull", rt.scaled_scene(200, max_depth=1), 72, 40, {"no_cull": 1}),
        ("scaled 200 no lrec", rt.scaled_scene(200, max_depth=1), 72, 40, {"no_cull": 2}),
        ("scaled 200 no packet", rt.scaled_scene(200, max_depth=1), 72, 40, {"no_cull": 4}),
        ("scaled 200 block kernel", rt.scaled_scene(200, max_depth=1), 72, 40, {"no_cull": 16}),
        ("scaled 6000 stream block kernel", rt.scaled_scene(6000, max_depth=0), 48, 28, {"no_cull": 16}),
        ("scaled 200 strict", rt.scaled_scene(200, max_depth=0), 48, 24, {"strict": True}),
        ("scaled 6000 stream", rt.scaled_scene(6000, max_depth=1), 48, 28, {}),
        ("scaled 6000 stream nocull", rt.scaled_scene(6000, max_depth=0), 48, 28, {"no_cull": 1}),
        ("scaled 6000 strict stream", rt.scaled_scene(6000, max_depth=0), 24, 12, {"strict": True})]:
    img, st = r.render(sc, cam, w, h, want_stats=True, **kw)
    tot += int(img.sum())
    print(name, img.shape, int(st["rays"]), int(st["cull_tests"]))
batch = r.render(rt.scaled_scene(100, max_depth=1), rt.turntable_cameras(3), 40, 24, spp=2)
print("batch", batch.shape, int(batch.sum()) + tot)
r.close()
