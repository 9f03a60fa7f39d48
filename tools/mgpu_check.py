"""Multi-GPU check, one process per GPU (torchrun): the row-band frame assembled by (a) the NCCL
gather and (b) the fused peer-store gather must be byte-identical to the single-GPU frame."""
import json, os, sys
import numpy as np
import torch
import torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import openglraytracer_b200 as rt
from openglraytracer_b200.multigpu import BandRenderer, PeerFrame, balanced_row_bands, probe_row_cost, row_band

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
tr = rt.Raytracer(local)
out = {}
for name, sc, w, h, spp in [("builtin_d4_1080p", rt.builtin_scene(0.0, 4), 1920, 1080, 1),
                            ("builtin_d4_4k_spp4", rt.builtin_scene(0.0, 4), 3840, 2160, 4),
                            ("scaled1k_d1_999x333", rt.scaled_scene(1024, max_depth=1), 999, 333, 1)]:
    cam = rt.builtin_camera(0.0)
    tr.upload_scene(sc)
    ref = tr.render(None, cam, w, h, spp=spp) if rank == 0 else None
    bands = balanced_row_bands(probe_row_cost(tr, None, cam, w, h), world)
    br = BandRenderer(tr, w, h, rank, world, dev, bands=bands)
    a = br.render_and_gather(None, cam, spp=spp)
    pf = PeerFrame(tr, w, h, rank, world, dev, bands=bands)
    pe = PeerFrame(tr, w, h, rank, world, dev)          # equal-height bands, for the imbalance figure
    b = pf.render(None, cam, spp=spp)
    torch.cuda.synchronize()
    if rank == 0:
        out[name] = {"nccl_gather_identical": bool(np.array_equal(a.cpu().numpy(), ref)),
                     "peer_store_identical": bool(np.array_equal(b.cpu().numpy(), ref))}
    # timing: 20 frames each way
    res = {}
    for label, fn in (("nccl_gather", lambda: br.render_and_gather(None, cam, spp=spp)), ("peer_store", lambda: pf.render(None, cam, spp=spp)),
                      ("peer_store_equal_height", lambda: pe.render(None, cam, spp=spp))):
        for _ in range(3):
            fn()
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / 20], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        res[label + "_ms"] = float(t.item())
    if rank == 0:
        out[name].update(res)
        out[name]["bands"] = str(bands)
    pf.close()
    pe.close()
if rank == 0:
    print(json.dumps({"world": world, **out}))
    assert all(v["nccl_gather_identical"] and v["peer_store_identical"] for v in out.values())
tr.close()
dist.barrier()
dist.destroy_process_group()
