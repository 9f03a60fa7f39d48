"""Weak-mode e2e under torchrun: every rank renders the default workload into its own page-locked frame, with the
kernel storing over PCIe (store) or a device frame + copy engine (copy).  Max over ranks, per mode."""
import os, sys, time, json
import numpy as np
import torch
import torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import openglraytracer_b200 as rt

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
r = rt.Raytracer(local)
sc, cam = rt.builtin_scene(0.0, 4), rt.builtin_camera(0.0)
w, h = 1920, 1080
pin = rt.PinnedImage((h, w, 4))
out = {}
for mode, label in ((2, "copy"), (1, "store"), (0, "auto")):
    for _ in range(10):
        r.render(sc, cam, w, h, out=pin.array, host_store=mode)
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    n = 100
    for _ in range(n):
        r.render(sc, cam, w, h, out=pin.array, host_store=mode)
    torch.cuda.synchronize()
    t = torch.tensor([(time.perf_counter() - t0) / n * 1e3], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    out[label] = round(float(t.item()), 4)
if rank == 0:
    print(json.dumps({"world": world, "e2e_ms_max_over_ranks": out}))
pin.free()
dist.destroy_process_group()
