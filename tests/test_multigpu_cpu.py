"""world_size-2 gloo test of the N>1 host logic on CPU: contiguous row bands,
padded gather on rank 0, assembled frame byte-identical to the single-rank frame.
The band pixels come from the oracle here (no GPU in this container); on the GPU
box tests/test_gpu_parity.py runs the same assembly with the CUDA kernel."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from openglraytracer_b200.multigpu import frame_block, frame_stride_set, gather_bands, row_band


def test_row_bands_partition_rows_exactly():
    for h in (1, 7, 270, 720, 1080, 2160, 2161):
        for g in (1, 2, 3, 4, 8):
            bands = [row_band(r, g, h) for r in range(g)]
            assert bands[0][0] == 0 and bands[-1][1] == h
            assert all(bands[i][1] == bands[i + 1][0] for i in range(g - 1))
            rows = [b - a for a, b in bands]
            assert max(rows) - min(rows) <= 1
    assert row_band(3, 8, 2160) == (810, 1080)       # 270-row bands at 4K / 8 GPUs
    assert [frame_block(r, 8, 256) for r in (0, 7)] == [(0, 32), (224, 256)]
    assert sorted(k for r in range(8) for k in frame_stride_set(r, 8, 250)) == list(range(250))
    assert frame_stride_set(3, 8, 256)[:3] == [3, 11, 19]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, w, h, out_path):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import oracle_binding as ob
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sc, cam = ob.builtin_scene(0.7, 2), ob.builtin_camera(0.7)
    y0, y1 = row_band(rank, world, h)
    full_canvas = ob.render(sc, cam, w, h, y0=y0, y1=y1, threads=2)     # absolute pixel coords (glsl:377-378)
    band = torch.from_numpy(np.ascontiguousarray(full_canvas[y0:y1]))
    frame = gather_bands(band, h, w)
    if rank == 0:
        np.save(out_path, frame.numpy())
    else:
        assert frame is None
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,h", [(2, 45), (3, 46)])
def test_band_gather_is_byte_identical_to_single_rank(tmp_path, world, h):
    import oracle_binding as ob
    w = 80
    out = str(tmp_path / "frame.npy")
    mp.spawn(_worker, args=(world, _free_port(), w, h, out), nprocs=world, join=True)
    ref = ob.render(ob.builtin_scene(0.7, 2), ob.builtin_camera(0.7), w, h)
    assert np.array_equal(np.load(out), ref)


def test_cost_balanced_bands_partition_and_balance():
    from openglraytracer_b200.multigpu import balanced_row_bands
    rng = np.random.default_rng(3)
    for h in (8, 45, 270, 1080, 2160):
        cost = rng.uniform(0.5, 1.5, h)
        cost[: h // 3] *= 4.0                       # floor / glass rows cost more rays than wall rows
        for g in (1, 2, 3, 4, 8):
            bands = balanced_row_bands(cost, g)
            assert bands[0][0] == 0 and bands[-1][1] == h and len(bands) == g
            assert all(bands[i][1] == bands[i + 1][0] for i in range(g - 1))
            assert all(b > a for a, b in bands)
            if h >= 270:
                per = [cost[a:b].sum() for a, b in bands]
                assert max(per) <= 1.08 * (cost.sum() / g) + 4 * cost.max()
                assert all(a % 4 == 0 for a, _ in bands)
    assert balanced_row_bands(np.zeros(16), 4) == [(0, 4), (4, 8), (8, 12), (12, 16)]


def _worker_uneven(rank, world, port, w, h, bands, out_path):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import oracle_binding as ob
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sc, cam = ob.builtin_scene(0.2, 1), ob.builtin_camera(0.2)
    y0, y1 = bands[rank]
    canvas = ob.render(sc, cam, w, h, y0=y0, y1=y1, threads=2)
    frame = gather_bands(torch.from_numpy(np.ascontiguousarray(canvas[y0:y1])), h, w, bands=bands)
    if rank == 0:
        np.save(out_path, frame.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_uneven_band_gather_is_byte_identical(tmp_path):
    import oracle_binding as ob
    w, h, bands = 64, 40, [(0, 8), (8, 40)]
    out = str(tmp_path / "frame.npy")
    mp.spawn(_worker_uneven, args=(2, _free_port(), w, h, bands, out), nprocs=2, join=True)
    assert np.array_equal(np.load(out), ob.render(ob.builtin_scene(0.2, 1), ob.builtin_camera(0.2), w, h))


def test_feedback_refinement_converges_on_a_synthetic_cost():
    from openglraytracer_b200.multigpu import refine_row_bands, row_band
    h, g = 2160, 8
    y = np.arange(h)
    true_cost = 1.0 + 4.0 * np.exp(-((y - 500) / 300.0) ** 2) + (y < 200) * 2.0     # rays per row, say

    def times(bs):
        return [float(true_cost[a:b].sum()) for a, b in bs]

    bands = [row_band(r, g, h) for r in range(g)]
    imbalance0 = max(times(bands)) / (true_cost.sum() / g)
    for _ in range(4):
        bands = refine_row_bands(bands, times(bands), h)
        assert bands[0][0] == 0 and bands[-1][1] == h and all(b > a for a, b in bands)
        assert all(bands[i][1] == bands[i + 1][0] for i in range(g - 1))
    imbalance = max(times(bands)) / (true_cost.sum() / g)
    assert imbalance0 > 1.5 and imbalance < 1.05


def test_numa_binding_is_best_effort():
    """bind_to_gpu_numa: without a usable NVML (this container) it leaves the process where it is and says so."""
    import os
    from openglraytracer_b200.multigpu import bind_to_gpu_numa
    before = os.sched_getaffinity(0)
    got = bind_to_gpu_numa(0)
    after = os.sched_getaffinity(0)
    assert got is None and after == before or (got is not None and set(got) == after and after <= before)
