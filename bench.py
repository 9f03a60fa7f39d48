#!/usr/bin/env python
"""bench.py — Mrays/s of the ray-scene kernel (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload NAME]

A "step" is one pass of the hot path over one frame of synthetic input: the
shader's built-in scene at 1920x1080, 1 sample/pixel, bounce depth 4
(BASELINE.json configs[1]); a "ray" is one get_closest_collision call of the
reference (glsl:738).  One JSON line is printed by rank 0.

  value        rays/s of the whole job, scene and camera already resident in HBM,
               framebuffer left in HBM; CUDA events on the launching stream, per
               step, L2 flushed between steps; max over ranks.
  e2e          same metric through the public call a user makes
               (Raytracer.render with HOST scene structs and a pinned HOST
               framebuffer): host prologue + H2D + kernel + D2H inside the timed region.
  roofline     FP32 CUDA-core roofline of the render kernel (north_star: no
               tensor cores, HBM irrelevant): algorithmic flops of the tests
               actually executed / kernel time, against the FP32 FMA peak
               measured in this run.
  cpu_baseline the oracle (CPU restatement of the shader, as written) timed on a
               bounded pixel sample of the same workload on the host cores.
  N > 1        every rank renders its own frame per step (frames shard with no
               data-path collective: "weak"); the row-band + NCCL-gather frame
               mode is timed as well and reported under "bands_gather".
  --impl reference   times the reference's CPU implementation of the path.  The
               path exists only as GLSL and no GL stack exists on this box
               ("unavailable on this box"), so this is the oracle port.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (scene kind, n_objects, width, height, spp, depth)
    "builtin_1080p_d4": ("builtin", 5, 1920, 1080, 1, 4),
    "builtin_1080p_d0": ("builtin", 5, 1920, 1080, 1, 0),
    "builtin_720p_d0": ("builtin", 5, 1280, 720, 1, 0),
    "builtin_4k_spp16_d4": ("builtin", 5, 3840, 2160, 16, 4),
    "scaled_1k_1080p_d0": ("scaled", 1024, 1920, 1080, 1, 0),
    "scaled_4k_1080p_d0": ("scaled", 4096, 1920, 1080, 1, 0),
    "scaled_16k_1080p_d0": ("scaled", 16384, 1920, 1080, 1, 0),
    "scaled_64k_1080p_d0": ("scaled", 65536, 1920, 1080, 1, 0),
    "scaled_1k_1080p_d2": ("scaled", 1024, 1920, 1080, 1, 2),
    "scaled_16k_1080p_d2": ("scaled", 16384, 1920, 1080, 1, 2),
    # BASELINE config 5: 256 turntable poses at 1024x1024, whole frames sharded over the ranks
    "batch256_1024_d4": ("batch", 5, 1024, 1024, 1, 4),
}
BATCH_FRAMES = 256
# dram__bytes_read.sum + dram__bytes_write.sum per launch of the timed kernel, from the committed
# `ncu --set full` captures (profiles/r1_ncu_full_*.txt); the 8.3 MB frame mostly stays in the 126 MB L2.
NCU_DRAM_BYTES = {"builtin_1080p_d4": 58368 + 1266688, "scaled_1k_1080p_d0": 8672512 + 167796736,
                  "scaled_64k_1080p_d0": 11898368 + 177194752}
# smsp__issue_active.avg.pct_of_peak_sustained_active of the timed kernel from the same captures: the share of issue
# slots in use (an FP32-issue-bound kernel's real ceiling; `frac` counts algorithmic flops only)
NCU_ISSUE_ACTIVE = {"builtin_1080p_d4": 0.830, "scaled_1k_1080p_d0": 0.451, "scaled_64k_1080p_d0": 0.526}
DEFAULT_WORKLOAD = "builtin_1080p_d4"
METRIC = "Mrays/s (rays x bounces)"
UNIT = "Mrays/s"

# Algorithmic FP32 flops per unit of work (add/sub/mul/min/max/cmp/div/sqrt = 1, FMA = 2;
# ray-dependent work only): box slab test 59, sphere test 41, closest-hit resolve 45, one shaded
# light 95, primary-ray generation 76.
F_BOX, F_SPH, F_HIT, F_LIGHT, F_PRIMARY = 59.0, 41.0, 45.0, 95.0, 76.0
# Large scenes run the slab test behind a bounding-sphere test and the sphere roots behind the
# discriminant: one such test is oc 3 + |oc|^2-r^2 6 (shared by the rays of one origin)
# + d.oc 5 + discriminant 3 + compare 1, counted as 12 per ray; a sphere test that goes on to its roots
# then costs the remaining 41 - 17 = 24.
F_CULL, F_SPH_ROOTS = 12.0, 24.0
# Packet (cone) test of one object against one cone, one object per lane:
# v = c - apex 3, h = v.axis 5, v x axis 9, p^2 5, rhs 2, far 1, three products 3, h^2 + p^2 2, compares 3.
F_PACKET = 26.0   # v 3, h 5, |v|^2 5, p^2 4, rhs 2, far 1, three products, three compares


def algorithmic_flops(st, samples):
    """Flops of the tests actually EXECUTED (skipped tests are never credited)."""
    f_sph = F_SPH_ROOTS if st["cull_tests"] else F_SPH
    return (F_BOX * st["box_tests"] + f_sph * st["sphere_tests"] + F_CULL * st["cull_tests"] +
            F_PACKET * st.get("packet_tests", 0) + F_HIT * st["hit_nodes"] + F_LIGHT * st["shadow_rays"] + F_PRIMARY * samples)


def make_scene(mod, kind, n, depth):
    if kind in ("builtin", "batch"):
        return mod.builtin_scene(0.0, depth)
    sc = mod.scaled_scene(n)
    sc.max_depth = depth
    return sc


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def stop(self, t0=None, t1=None):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.proc.terminate()
        sm, smax, reasons = [], [], set()
        for ts, line in self.rows:
            if t0 is not None and not (t0 - 0.06 <= ts <= t1 + 0.12):
                continue
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                smax.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------
# CPU legs (the ONLY places bench.py touches oracle/: cpu_baseline and --impl reference)
# ----------------------------------------------------------------------------
def _oracle():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_binding as ob
    return ob


def strided_pixels(width, height, n_pixels, seed=1):
    """A bounded, uniformly spread sample of the frame's pixels (same workload mix as the frame)."""
    total = width * height
    n = min(n_pixels, total)
    rng = np.random.default_rng(seed)
    idx = np.sort(rng.choice(total, size=n, replace=False))
    return np.stack([idx % width, idx // width], axis=1).astype(np.int32)


def cpu_time_sample(ob, wl, target_seconds, as_written=True):
    """Time the oracle on a pixel sample sized for about target_seconds; returns (Mrays/s, description, cores)."""
    kind, n, w, h, spp, depth = wl
    scene = make_scene(ob, kind, n, depth) if kind in ("builtin", "batch") else None
    if scene is None:
        import openglraytracer_b200 as rt   # scene GENERATOR only (host data); nothing is rendered by the product here
        scene = make_scene(rt, kind, n, depth)
    cam = ob.builtin_camera(0.0)
    cores = ob.lib().oracle_num_threads()
    probe = strided_pixels(w, h, 2048)
    t0 = time.perf_counter()
    _, st = ob.render(scene, cam, w, h, spp=spp, pixels=probe, as_written=as_written, want_stats=True)
    dt = max(time.perf_counter() - t0, 1e-4)
    n_pix = int(min(w * h, max(2048, len(probe) * target_seconds / dt)))
    pix = strided_pixels(w, h, n_pix, seed=2)
    t0 = time.perf_counter()
    _, st = ob.render(scene, cam, w, h, spp=spp, pixels=pix, as_written=as_written, want_stats=True)
    dt = time.perf_counter() - t0
    desc = "%d of %d pixels (seeded uniform subset) of one frame, %.2f s" % (len(pix), w * h, dt)
    return st["rays"] / dt / 1e6, desc, cores, st["rays"], dt


def run_reference(args, wl_name, wl):
    """--impl reference: the reference's own CPU implementation of the path.  The
    path exists only as a GLSL compute shader; OpenGL / Mesa llvmpipe are
    unavailable on this box, so the CPU arm is the oracle port executing the
    shader's arithmetic as written, on all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    ob = _oracle()
    kind, n, w, h, spp, depth = wl
    per_step = 20.0 / max(args.steps + args.warmup, 1)
    per_step = min(max(per_step, 0.5), 6.0)
    for _ in range(args.warmup):
        cpu_time_sample(ob, wl, per_step)
    rays = 0
    secs = 0.0
    desc = ""
    for _ in range(args.steps):
        v, desc, cores, r, dt = cpu_time_sample(ob, wl, per_step)
        rays += r
        secs += dt
    value = rays / secs / 1e6
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / max(args.steps, 1),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl_name, "scene": kind, "objects": n, "width": w, "height": h, "spp": spp,
                   "max_depth": depth, "note": "GL/llvmpipe unavailable on this box; oracle port of the shader, as written"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc + " per step"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------
def run_gpu(args, wl_name, wl):
    import torch
    import torch.distributed as dist
    import openglraytracer_b200 as rt
    from openglraytracer_b200.multigpu import (BandRenderer, PeerFrame, balanced_row_bands, frame_stride_set,
                                                 probe_row_cost, row_band)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa_cpus = None
    if world > 1:
        from openglraytracer_b200.multigpu import bind_to_gpu_numa
        numa_cpus = bind_to_gpu_numa(local)         # before any page-locked allocation
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    kind, n_obj, w, h, spp, depth = wl
    scene = make_scene(rt, kind, n_obj, depth)
    cam = rt.builtin_camera(0.0)
    batch = kind == "batch"
    nfr = 1
    cams = cam
    if batch:   # whole frames per rank: poses k = rank (mod world) of the 256-pose turntable (glsl:343-358 orbit)
        mine = frame_stride_set(rank, world, BATCH_FRAMES)
        poses = rt.turntable_cameras(BATCH_FRAMES)
        nfr = len(mine)
        cams = (rt.Camera * nfr)(*[poses[k] for k in mine])
    tracer = rt.Raytracer(local)
    # a real (non-NULL) stream: the C ABI treats a NULL stream as "the context's own stream"
    stream = torch.cuda.Stream(dev)
    torch.cuda.set_stream(stream)
    frame_dev = torch.empty((nfr, h, w, 4), dtype=torch.uint8, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)       # > 126 MB L2

    # counters from the instrumented twin (untimed); the timed kernel is the plain one
    tracer.upload_scene(scene)
    _, st = tracer.render(None, cams, w, h, spp=spp, out_device_ptr=frame_dev.data_ptr(), want_stats=True)
    rays_per_frame = int(st["rays"])            # per launch of this rank (nfr frames in batch mode)
    flops_per_frame = algorithmic_flops(st, nfr * w * h * spp)
    peak_tf, _ = rt.fp32_peak_tflops(local, 8192)

    def step():
        tracer.render(None, cams, w, h, spp=spp, out_device_ptr=frame_dev.data_ptr(), async_=True,
                      stream=stream.cuda_stream)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for _ in range(max(args.warmup, 3)):
        step()
        flush.zero_()
    barrier()
    launches0 = rt.kernel_launch_count()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    t_region0 = time.perf_counter()
    for a, b in evs:
        flush.zero_()                 # L2 flush between timed iterations (outside the event pair)
        a.record(stream)
        step()
        b.record(stream)
    barrier()
    t_region1 = time.perf_counter()
    launches = rt.kernel_launch_count() - launches0
    step_ms = [a.elapsed_time(b) for a, b in evs]
    total_ms = float(sum(step_ms))
    if world > 1:
        t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
    ms_per_step = total_ms / args.steps
    rays_job = world * rays_per_frame           # frame mode: every rank renders its own copy of the frame (weak)
    if batch and world > 1:                     # batch mode: the 256 poses are split over the ranks (strong)
        t = torch.tensor([rays_per_frame], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        rays_job = int(t.item())
    value = rays_job / (ms_per_step * 1e-3) / 1e6

    # e2e: the public call with host structs and a pinned host framebuffer
    pin = rt.PinnedImage((nfr, h, w, 4) if batch else (h, w, 4))
    scene_bytes = 132 * scene.objects_count + 60 * scene.scene.lights_count + 40 * nfr
    for _ in range(3):
        tracer.render(scene, cams, w, h, spp=spp, out=pin.array)
    barrier()
    e2e_steps = max(3, min(args.steps, 50))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        tracer.render(scene, cams, w, h, spp=spp, out=pin.array)
    barrier()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e_value = rays_job * e2e_steps / e2e_s / 1e6
    checksum = int(pin.array[..., ::97, ::89, :].astype(np.uint64).sum())

    # frame mode at N > 1: one frame split into contiguous row bands, NCCL gather on rank 0
    bands = None
    if world > 1 and not batch:
        # contiguous row bands of equal COST (per-row ray counts from a 72-strip probe, same on every rank)
        cost_bands = balanced_row_bands(probe_row_cost(tracer, None, cam, w, h), world)
        torch.cuda.synchronize(dev)
        br = BandRenderer(tracer, w, h, rank, world, dev, bands=cost_bands)
        for _ in range(3):
            br.render_and_gather(None, cam, spp=spp)
        barrier()
        n_b = max(3, min(args.steps, 50))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(n_b):
            full = br.render_and_gather(None, cam, spp=spp)
        e1.record(stream)
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_b = float(t.item()) / n_b
        same = True
        if rank == 0:
            same = bool(np.array_equal(full.cpu().numpy(), pin.array))
        bands = {"ms_per_frame": ms_b, "value": rays_per_frame / (ms_b * 1e-3) / 1e6, "unit": UNIT,
                 "scaling": "strong", "collective": "nccl gather of %d row bands to rank 0" % world,
                 "bands": "contiguous, cost-balanced: " + str(cost_bands),
                 "byte_identical_to_single_gpu": same}
        # fused gather: every rank's kernel stores its band straight into rank 0's frame (CUDA IPC, NVLink)
        try:
            for label, bset in (("peer_store", cost_bands), ("peer_store_feedback_bands", None),
                                ("peer_store_equal_height_bands", [row_band(r, world, h) for r in range(world)])):
                pf = PeerFrame(tracer, w, h, rank, world, dev, bands=bset or cost_bands)
                if bset is None:     # three feedback steps on the measured per-band times
                    bset = pf.autotune(None, cam, cost_bands, iters=3, spp=spp)
                for _ in range(3):
                    pf.render(None, cam, spp=spp)
                barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                for _ in range(n_b):
                    full = pf.render(None, cam, spp=spp)
                e1.record(stream)
                barrier()
                t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ms_p = float(t.item()) / n_b
                same_p = bool(np.array_equal(full.cpu().numpy(), pin.array)) if rank == 0 else True
                pf.close()
                bands[label] = {"ms_per_frame": ms_p, "value": rays_per_frame / (ms_p * 1e-3) / 1e6,
                                "collective": "none on the data path: kernels store into rank 0's frame over "
                                              "NVLink (CUDA IPC mapping), one 4-byte all-reduce as completion flag",
                                "bands": str(bset), "byte_identical_to_single_gpu": same_p}
        except Exception as exc:   # IPC refused (container policy): the NCCL gather above stands
            bands["peer_store"] = {"unavailable": repr(exc)[:200]}

    clocks = sampler.stop(t_region0, t_region1) if rank == 0 else None
    if rank == 0:
        kernel_ms = float(np.mean(step_ms))
        achieved = flops_per_frame / (kernel_ms * 1e-3) / 1e12
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong" if batch else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl_name, "scene": kind, "objects": n_obj, "width": w, "height": h, "spp": spp,
                       "max_depth": depth, "rays_per_frame": rays_per_frame // nfr,
                       "rays_traversed_per_frame": (rays_per_frame - int(st["rays_not_traversed"])) // nfr,
                       "ray_unit": "one get_closest_collision call of the reference (glsl:738); shadow rays whose outcome "
                                   "cannot change the pixel are counted but not traversed",
                       "frames_per_step": BATCH_FRAMES if batch else world,
                       "parallelism": "frames sharded over %d GPU(s), no data-path collective" % world,
                       "host_binding": ("each rank on the %d cores next to its GPU (NVML affinity)" % len(numa_cpus)) if numa_cpus else "launcher default",
                       "l2": "flushed between timed steps (256 MiB write)", "kernel": "fast (fmad on, no fast-math)"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": scene_bytes,
                    "d2h_bytes_per_step": nfr * w * h * 4, "steps": e2e_steps, "ms_per_step": 1e3 * e2e_s / e2e_steps,
                    "checksum": checksum},
            "gpu_launches": int(launches),
            "roofline": {"bound": "fp32", "achieved": achieved, "peak": peak_tf, "unit": "TFLOP/s",
                         "frac": achieved / peak_tf, "traffic": NCU_DRAM_BYTES.get(wl_name),
                         "issue_slots_active_ncu": NCU_ISSUE_ACTIVE.get(wl_name),
                         "traffic_note": "DRAM bytes per launch (ncu --set full, profiles/); algorithmic bytes = the "
                                         "%d-byte frame + the scene records: FP32-issue bound, not HBM; the large-scene "
                                         "kernels also write back the local-memory frames of their resident threads (1.8 KB x "
                                         "75 776 threads) once per launch" % (nfr * w * h * 4),
                         "peak_source": "FP32 FMA-chain microbenchmark in this run (MEASURED_PEAKS.json has no FP32 "
                                        "figure); nominal 148 SM x 128 x 2 x 1.965 GHz = 74.4",
                         "algorithmic_flops_per_launch": flops_per_frame, "kernel_ms": kernel_ms,
                         "box_tests": int(st["box_tests"]), "sphere_tests": int(st["sphere_tests"]),
                         "bounding_sphere_tests": int(st["cull_tests"]),
                         "packet_cone_tests": int(st.get("packet_tests", 0)),
                         "lane_use": st["nodes"] / max(1, 32 * st["warp_iters"]),
                         "flops_per_unit": "slab test 59, sphere 41 (24 after its discriminant), bounding-sphere / "
                                           "discriminant test 12, packet cone test 26 per (object, cone), hit resolve 45, "
                                           "shaded light 95, primary ray 76; executed tests only"},
            "clocks": clocks,
        }
        if bands:
            line["bands_gather"] = bands
        if world == 1 and not args.no_cpu_baseline:
            ob = _oracle()
            v, desc, cores, _, _ = cpu_time_sample(ob, wl, args.cpu_seconds)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc}
        print(json.dumps(line), flush=True)
    pin.free()
    tracer.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, args.workload, wl)
    else:
        run_gpu(args, args.workload, wl)


if __name__ == "__main__":
    main()
