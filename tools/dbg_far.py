"""Debug helper: far/tiny object scene, strict vs fast vs oracle."""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import openglraytracer_b200 as rt
import oracle_binding as ob

def scene(variant):
    base = rt.builtin_scene(0.0, 1)
    n = 64
    objs = (rt.Object * n)()
    rng = np.random.default_rng(7)
    for i in range(n):
        o = objs[i]
        o.sphere_radius = -1.0
        o.material = base.objects[1 + i % 4].material
        far = 2000.0 if i % 3 == 0 else 8.0
        ctr = rng.uniform(-1, 1, 3) * far
        o.position[:] = [float(x) for x in ctr]
        o.angles[:] = [float(x) for x in rng.uniform(0, 360, 3)]
        if i % 7 == 6:
            o.sphere_radius = float(rng.uniform(0.2, 300.0 if far > 100 else 2.0))
        else:
            hx = rng.uniform(0.001, 0.01, 3) if i % 5 == 0 else rng.uniform(0.2, 200.0 if far > 100 else 2.0, 3)
            o.box_mins[:] = [float(-x) for x in hx]
            o.box_maxs[:] = [float(x) for x in hx]
    objs[0] = base.objects[0]
    big = 3000.0 if variant & 1 else 11.0
    objs[0].box_mins[:] = [-big] * 3
    objs[0].box_maxs[:] = [big] * 3
    if variant & 2:
        for i in range(1, n):
            if abs(objs[i].position[0]) > 100 or abs(objs[i].position[1]) > 100 or abs(objs[i].position[2]) > 100:
                objs[i].sphere_radius = -1.0
                objs[i].box_mins[:] = [0, 0, 0]; objs[i].box_maxs[:] = [0, 0, 0]
    return rt.SceneData(objs, base.lights, 2)

r = rt.Raytracer()
cam = rt.builtin_camera(0.0)
for variant in range(4):
    sc = scene(variant)
    ref = ob.render(sc, cam, 320, 180)
    for strict in (True, False):
        img = r.render(sc, cam, 320, 180, strict=strict)
        d = np.abs(img.astype(int) - ref.astype(int)).max(axis=2)
        print("variant", variant, "strict" if strict else "fast", "over1", int((d > 1).sum()), "max", d.max())
    ys, xs = np.nonzero(d > 1)
    for y, x in list(zip(ys, xs))[:5]:
        print("   ", x, y, img[y, x], ref[y, x])
