"""Renders one scaled-scene workload a few times (ncu target)."""
import sys, os, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import openglraytracer_b200 as rt

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
w, h = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (1920, 1080)
depth = int(sys.argv[4]) if len(sys.argv) > 4 else 0
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 3
no_cull = {"nocull": 1, "nopacket": 4, "nopacket_nolrec": 6, "noblock": 8, "block": 16, "nofan": 32, "nogroup": 64, "block_nogroup": 80, "noblock_nogroup": 72}.get(sys.argv[6], 0) if len(sys.argv) > 6 else 0
sc = rt.scaled_scene(n, max_depth=depth) if n > 5 else rt.builtin_scene(0.0, depth)
cam = rt.builtin_camera(0.0)
r = rt.Raytracer()
r.upload_scene(sc)
pk, _ = rt.fp32_peak_tflops(iters=4096)
for i in range(reps):
    img, st = r.render(None, cam, w, h, want_stats=True, no_cull=no_cull)
    fl = 59.0 * st["box_tests"] + (24.0 if st["cull_tests"] else 41.0) * st["sphere_tests"] + 12.0 * st["cull_tests"] + 26.0 * st["packet_tests"]
    print(json.dumps(dict(n=n, w=w, h=h, depth=depth, kernel_ms=st["kernel_ms"], tflops=fl / st["kernel_ms"] / 1e9,
                          frac=fl / st["kernel_ms"] / 1e9 / pk, peak=pk, mrays=st["rays"] / st["kernel_ms"] / 1e3,
                          box_tests=st["box_tests"], sph_tests=st["sphere_tests"], cull_tests=st["cull_tests"], packet_tests=st["packet_tests"], rays=st["rays"], nodes=st["nodes"], warp_iters=st["warp_iters"], group_passes=st.get("group_passes", 0), loop_passes=st.get("loop_passes", 0), candidates=st.get("candidates", 0), shadow_group_passes=st.get("shadow_group_passes", 0),
                          lane_use=st["nodes"] / max(1, 32 * st["warp_iters"]), no_cull=no_cull)))
