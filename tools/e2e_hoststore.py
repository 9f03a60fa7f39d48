"""e2e frame time through the public call with a pinned host frame: kernel stores over PCIe vs D2H copy."""
import sys, os, time, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import openglraytracer_b200 as rt
r = rt.Raytracer()
cam = rt.builtin_camera(0.0)
for name, sc, w, h, spp in [("builtin d0 1080p", rt.builtin_scene(0.0, 0), 1920, 1080, 1), ("builtin d1 1080p", rt.builtin_scene(0.0, 1), 1920, 1080, 1),
                            ("builtin d4 1080p", rt.builtin_scene(0.0, 4), 1920, 1080, 1), ("builtin d4 4k", rt.builtin_scene(0.0, 4), 3840, 2160, 1),
                            ("scaled 1k d0 1080p", rt.scaled_scene(1024), 1920, 1080, 1)]:
    pin = rt.PinnedImage((h, w, 4))
    out = {}
    ref = None
    for mode, label in ((2, "copy"), (1, "store"), (0, "auto")):
        for _ in range(5):
            r.render(sc, cam, w, h, spp=spp, out=pin.array, host_store=mode)
        t0 = time.perf_counter()
        n = 50
        for _ in range(n):
            r.render(sc, cam, w, h, spp=spp, out=pin.array, host_store=mode)
        out[label] = round((time.perf_counter() - t0) / n * 1e3, 4)
        img = pin.array.copy()
        if ref is None: ref = img
        out[label + "_same"] = bool(np.array_equal(img, ref))
    print(name, json.dumps(out))
    pin.free()
