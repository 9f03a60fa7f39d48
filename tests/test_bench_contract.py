"""bench.py's contract, as far as it can be checked without a GPU: the reference arm (`--impl reference`, the CPU
port of the shader timed on the host cores) prints ONE JSON line with the keys the driver reads, under torchrun
only rank 0 prints; the product arm fails loudly when there is no CUDA device (no CPU fallback)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, env=None, timeout=300):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, capture_output=True, text=True,
                          timeout=timeout, env=e, cwd=ROOT)


def test_reference_arm_prints_the_contract_line():
    r = _run(["--impl", "reference", "--steps", "1", "--warmup", "0", "--workload", "builtin_720p_d0"])
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["higher_is_better"] is True and d["vs_baseline"] is None
    assert d["metric"].startswith("Mrays/s") and d["unit"] == "Mrays/s" and d["value"] > 0
    assert d["n_gpus"] == 1 and d["steps"] == 1 and d["warmup"] == 0 and d["ms_per_step"] > 0
    assert d["config"]["workload"] == "builtin_720p_d0" and d["dtype"] == "f32" and d["data"] == "synthetic"
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["gpu_launches"] == 0


def test_reference_arm_other_ranks_stay_silent():
    r = _run(["--impl", "reference", "--steps", "1", "--warmup", "0", "--gpus", "2"], env={"RANK": "1", "WORLD_SIZE": "2"})
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_product_arm_has_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = _run(["--steps", "1", "--warmup", "0"])
    assert r.returncode != 0
    assert "no CUDA device" in (r.stderr + r.stdout) and "no CPU fallback" in (r.stderr + r.stdout)


def test_reference_arm_uses_all_host_threads_under_torchrun():
    """torch.distributed.run exports OMP_NUM_THREADS=1 to its workers; the CPU arm must still use every host
    thread (round 1's N > 1 reference lines ran on one thread) and describe the workload with the very dict the
    GPU arm prints, whatever --gpus says."""
    sys.path.insert(0, ROOT)
    import bench
    r = _run(["--impl", "reference", "--steps", "1", "--warmup", "0", "--gpus", "2", "--workload", "builtin_720p_d0"],
             env={"RANK": "0", "WORLD_SIZE": "2", "LOCAL_RANK": "0", "OMP_NUM_THREADS": "1"})
    assert r.returncode == 0, r.stderr[-2000:]
    d = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][0])
    assert d["cpu_baseline"]["cores"] == len(os.sched_getaffinity(0))
    assert d["config"] == bench.static_config("builtin_720p_d0", bench.WORKLOADS["builtin_720p_d0"], 2)
    if len(os.sched_getaffinity(0)) >= 4:
        # not a one-thread run: same ball park as the plain launch.  Two short CPU timings on a shared host: the
        # better of two runs each, and a wide margin (a one-thread run would sit at 1 / cores)
        def best(args, env=None):
            vals = []
            for _ in range(2):
                rr = _run(args, env=env)
                vals.append(json.loads([l for l in rr.stdout.splitlines() if l.startswith("{")][0])["value"])
            return max(vals)
        v1 = best(["--impl", "reference", "--steps", "1", "--warmup", "0", "--workload", "builtin_720p_d0"])
        vt = max(d["value"], best(["--impl", "reference", "--steps", "1", "--warmup", "0", "--gpus", "2", "--workload", "builtin_720p_d0"],
                                  env={"RANK": "0", "WORLD_SIZE": "2", "LOCAL_RANK": "0", "OMP_NUM_THREADS": "1"}))
        assert vt > 0.4 * v1


def test_roofline_flops_credit_executed_work_only():
    """bench.algorithmic_flops: 95 flop per SHADED light (not per shadow ray the reference would cast), 5 per
    traversed shadow ray, nothing for untraversed ones."""
    sys.path.insert(0, ROOT)
    import bench
    st = dict(box_tests=0, sphere_tests=0, cull_tests=0, packet_tests=0, hit_nodes=0, shaded_lights=0,
              shadow_rays=300, rays_not_traversed=100)
    assert bench.algorithmic_flops(st, 0) == 5.0 * 200
    st["shaded_lights"] = 120
    assert bench.algorithmic_flops(st, 0) == 5.0 * 200 + 95.0 * 120
    st["rays_not_traversed"] = 300
    st["shaded_lights"] = 0
    assert bench.algorithmic_flops(st, 10) == 76.0 * 10


def test_ncu_figures_are_tied_to_the_kernel_sources(tmp_path, monkeypatch):
    sys.path.insert(0, ROOT)
    import bench
    sha = bench.csrc_sha()
    prof = tmp_path / "profiles"
    prof.mkdir()
    monkeypatch.setattr(bench, "ROOT", str(tmp_path))
    monkeypatch.setattr(bench, "csrc_sha", lambda: sha)
    assert bench.ncu_capture("builtin_1080p_d4")[0] is None
    ent = {"csrc_sha": sha, "file": "profiles/x.txt", "dram_bytes_read": 10, "dram_bytes_write": 5, "issue_active": 0.8}
    (prof / "ncu_captures.json").write_text(json.dumps({"kernels": {"builtin_1080p_d4": ent}}))
    assert bench.ncu_capture("builtin_1080p_d4")[:2] == (15, 0.8)
    ent["csrc_sha"] = "0" * 16
    (prof / "ncu_captures.json").write_text(json.dumps({"kernels": {"builtin_1080p_d4": ent}}))
    t, i, why = bench.ncu_capture("builtin_1080p_d4")
    assert t is None and i is None and "stale" in why
