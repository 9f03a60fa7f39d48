"""Generates the golden fixtures from the CPU oracle (run in the build container).

The reference ships no golden vectors and its shader cannot run here (no GL), so
these pin the ORACLE against itself over time; the GPU parity tests compare the
CUDA kernel with these same files on the GPU box.
  golden.json        : sha256 of full oracle frames
  golden_images.npz  : small full RGBA frames (named t<time>_d<depth>_w<W>_h<H>)
  golden_scaled.npz  : small full frames of the scaled synthetic scene (BASELINE config 4; named
                       n<objects>_d<depth>_w<W>_h<H>); the scene comes from the library's host-side generator
                       (rt_scaled_scene, no GPU needed), golden.json keeps the sha256 of its object bytes
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_binding as ob  # noqa: E402

CASES = [  # time, depth, W, H, spp, seed
    (0.0, 0, 1280, 720, 1, 0x5EED0001), (0.0, 4, 640, 360, 1, 0x5EED0001), (2.5, 2, 640, 360, 1, 0x5EED0001),
    (3.7, 4, 320, 180, 1, 0x5EED0001), (0.0, 2, 160, 90, 16, 0x5EED0001), (11.0, 8, 96, 54, 1, 0x5EED0001),
    (0.0, 1, 33, 17, 1, 0x5EED0001),
]
SCALED = [(1024, 1, 96, 54), (4096, 0, 96, 54), (5000, 2, 64, 36), (20000, 0, 64, 36)]   # objects, depth, W, H
IMAGES = [(0.0, 0, 160, 90), (0.0, 4, 160, 90), (2.5, 2, 128, 72), (3.7, 4, 128, 72), (7.25, 6, 64, 36),
          (3.7, 15, 64, 36)]


def main():
    cases = []
    for t, d, w, h, spp, seed in CASES:
        img = ob.render(ob.builtin_scene(t, d), ob.builtin_camera(t), w, h, spp=spp, seed=seed)
        cases.append(dict(time=t, depth=d, width=w, height=h, spp=spp, seed=seed,
                          sha256=hashlib.sha256(img.tobytes()).hexdigest()))
    sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
    import openglraytracer_b200 as rt
    scaled, simgs = [], {}
    for n, d, w, h in SCALED:
        sc = rt.scaled_scene(n, max_depth=d)
        simgs["n%d_d%d_w%d_h%d" % (n, d, w, h)] = ob.render(sc, rt.builtin_camera(0.0), w, h)
        scaled.append(dict(objects=n, depth=d, width=w, height=h, scene_sha256=hashlib.sha256(bytes(sc.objects)).hexdigest()))
    np.savez_compressed(os.path.join(HERE, "golden_scaled.npz"), **simgs)
    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(dict(generator="tests/golden/make_golden.py", oracle="oracle/rt_oracle.cpp", cases=cases, scaled=scaled), f, indent=1)
    imgs = {}
    for t, d, w, h in IMAGES:
        imgs["t%g_d%d_w%d_h%d" % (t, d, w, h)] = ob.render(ob.builtin_scene(t, d), ob.builtin_camera(t), w, h)
    np.savez_compressed(os.path.join(HERE, "golden_images.npz"), **imgs)
    print("wrote", len(cases), "hashes,", len(imgs), "images and", len(simgs), "scaled-scene images")


if __name__ == "__main__":
    main()
