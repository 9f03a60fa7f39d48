import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import openglraytracer_b200 as rt
from test_gpu_parity import _far_scene
r = rt.Raytracer()
cam = rt.builtin_camera(0.0)
def null(o):
    o.sphere_radius = -1.0
    o.box_mins[:] = [0, 0, 0]; o.box_maxs[:] = [0, 0, 0]
for variant in ("all", "nospheres", "nearonly", "faronly", "notiny", "depth0"):
    sc = _far_scene(3000.0)
    for i in range(2, 64):
        o = sc.objects[i]
        far = max(abs(o.position[0]), abs(o.position[1]), abs(o.position[2])) > 50
        sph = o.sphere_radius != -1.0
        tiny = (not sph) and o.box_maxs[0] < 0.05
        if variant == "nospheres" and sph: null(o)
        if variant == "nearonly" and far: null(o)
        if variant == "faronly" and not far: null(o)
        if variant == "notiny" and tiny: null(o)
    if variant == "depth0": sc.max_depth = 0
    imgs = {k: r.render(sc, cam, 320, 180, no_cull=k) for k in (0, 1, 2)}
    d = np.abs(imgs[0].astype(int) - imgs[1].astype(int)).max(axis=2)
    ys, xs = np.nonzero(d)
    print(variant, "lrec vs plain: diff px", len(ys), "max", d.max(), list(zip(xs[:4].tolist(), ys[:4].tolist())),
          "generic vs plain", int((imgs[2] != imgs[1]).sum()))
