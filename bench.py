#!/usr/bin/env python
"""bench.py — Mrays/s of the ray-scene kernel (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload NAME]

A "step" is one pass of the hot path over one frame of synthetic input: the
shader's built-in scene at 1920x1080, 1 sample/pixel, bounce depth 4
(BASELINE.json configs[1]); a "ray" is one get_closest_collision call of the
reference (glsl:738).  One JSON line is printed by rank 0.

  value        rays/s of the whole job, scene and camera already resident in HBM,
               framebuffer left in HBM; CUDA events on the launching stream, per
               step, L2 flushed between steps; max over ranks.  Counts the rays the
               REFERENCE would trace; value_traversed counts only the rays the kernel
               actually traverses (shadow rays that cannot change the pixel are skipped).
  e2e          same metric through the public call a user makes
               (Raytracer.render with HOST scene structs and a pinned HOST
               framebuffer): host prologue + H2D + kernel + D2H inside the timed region.
  roofline     FP32 CUDA-core roofline of the render kernel (north_star: no
               tensor cores, HBM irrelevant): algorithmic flops of the work
               actually executed / kernel time, against the FP32 FMA peak
               measured in this run.
  cpu_baseline the oracle (CPU restatement of the shader, as written) timed on a
               bounded pixel sample of the same workload on the host cores.
  scaled       (N = 1) BASELINE config 4, the scaled synthetic scene at 1k and 64k
               objects, shipped kernel AND with every slab test executed (no_cull=1,
               the loop the north-star 60 % target is stated on).
  N > 1        every rank renders its own frame per step (frames shard with no
               data-path collective: "weak"); the one-frame modes are timed as extra
               keys: bands_gather (this workload), bands_gather_4k_spp16 (config 3)
               and batch256 (config 5).
  --impl reference   times the reference's CPU implementation of the path.  The
               path exists only as GLSL and no GL stack exists on this box
               ("unavailable on this box"), so this is the oracle port.
"""
import argparse
import glob
import hashlib
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
    "builtin_720p_d0": ("builtin", 5, 1280, 720, 1, 0),     # BASELINE config 1 (the reference's window, shader as shipped)
    "builtin_720p_d4": ("builtin", 5, 1280, 720, 1, 4),
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
DEFAULT_WORKLOAD = "builtin_1080p_d4"
METRIC = "Mrays/s (rays x bounces)"
UNIT = "Mrays/s"
SCALED_LEGS = ("scaled_1k_1080p_d0", "scaled_64k_1080p_d0", "scaled_1k_1080p_d2", "scaled_16k_1080p_d2")
SCALED_LEGS_SHIPPED_ONLY = ("scaled_1k_1080p_d2", "scaled_16k_1080p_d2")   # every slab test at depth 2 takes seconds per frame

# Algorithmic FP32 flops per unit of work (add/sub/mul/min/max/cmp/div/sqrt = 1, FMA = 2;
# ray-dependent work only): box slab test 59, sphere test 41, closest-hit resolve 45, one SHADED
# light 95 (glsl:822-833: only for a shadow ray that was traversed and found no occluder), a traversed
# shadow ray's set-up 5 (origin p + 0.01 n shared by the lights of a node, direction L - p), primary ray 76.
F_BOX, F_SPH, F_HIT, F_LIGHT, F_SHADOW_RAY, F_PRIMARY = 59.0, 41.0, 45.0, 95.0, 5.0, 76.0
# Large scenes run the slab test behind a bounding-sphere test and the sphere roots behind the
# discriminant: one such test is oc 3 + |oc|^2-r^2 6 (shared by the rays of one origin)
# + d.oc 5 + discriminant 3 + compare 1, counted as 12 per ray; a sphere test that goes on to its roots
# then costs the remaining 41 - 17 = 24.
F_CULL, F_SPH_ROOTS = 12.0, 24.0
# Packet (cone) test of one object against one cone, one object per lane:
F_PACKET = 26.0   # v 3, h 5, |v|^2 5, p^2 4, rhs 2, far 1, three products, three compares


def algorithmic_flops(st, samples):
    """Flops of the work actually EXECUTED.  Skipped tests, shadow rays that were never traversed and lights
    that were not shaded (occluded, zero-weight surface, unlit light) are not credited."""
    f_sph = F_SPH_ROOTS if st["cull_tests"] else F_SPH
    traversed_shadow = st["shadow_rays"] - st["rays_not_traversed"]
    return (F_BOX * st["box_tests"] + f_sph * st["sphere_tests"] + F_CULL * st["cull_tests"] +
            F_PACKET * st.get("packet_tests", 0) + F_HIT * st["hit_nodes"] + F_LIGHT * st["shaded_lights"] +
            F_SHADOW_RAY * traversed_shadow + F_PRIMARY * samples)


FLOPS_NOTE = ("slab test 59, sphere 41 (24 after its discriminant), bounding-sphere / discriminant test 12, packet cone "
              "test 26 per (object, cone), hit resolve 45, SHADED light 95 (counted where glsl:822-833 runs), traversed "
              "shadow ray 5, primary ray 76; executed work only -- untraversed shadow rays and occluded lights earn nothing")


def make_scene(mod, kind, n, depth):
    if kind in ("builtin", "batch"):
        return mod.builtin_scene(0.0, depth)
    sc = mod.scaled_scene(n)
    sc.max_depth = depth
    return sc


def static_config(wl_name, wl, world):
    """What both arms (GPU and --impl reference) say about the workload: identical dicts."""
    kind, n, w, h, spp, depth = wl
    batch = kind == "batch"
    return {"workload": wl_name, "scene": kind, "objects": n, "width": w, "height": h, "spp": spp, "max_depth": depth,
            "frames_per_step": BATCH_FRAMES if batch else world,
            "parallelism": "frames sharded over %d GPU(s), no data-path collective" % world,
            "ray_unit": "one get_closest_collision call of the reference (glsl:738)",
            "l2": "GPU arm: flushed between timed steps (256 MiB write)"}


# ----------------------------------------------------------------------------
# ncu figures (traffic, issue-slot use) come from a committed capture that names the kernel sources
# it was taken on; a capture of other sources is not reported
# ----------------------------------------------------------------------------
def csrc_sha():
    h = hashlib.sha256()
    d = os.path.join(ROOT, "openglraytracer_b200", "csrc")
    for f in sorted(glob.glob(os.path.join(d, "rt_kernel*")) + [os.path.join(d, "rt_cone.h"), os.path.join(d, "rt_device.h"),
                                                                    os.path.join(d, "rt_groups.cu")]):
        with open(f, "rb") as fh:
            h.update(os.path.basename(f).encode() + b"\0" + fh.read())
    return h.hexdigest()[:16]


def ncu_capture(wl_name):
    """(traffic bytes per launch, issue-slot share, note) from profiles/ncu_captures.json, or (None, None, why)."""
    path = os.path.join(ROOT, "profiles", "ncu_captures.json")
    try:
        with open(path) as f:
            cap = json.load(f)
    except Exception:
        return None, None, "no profiles/ncu_captures.json"
    e = cap.get("kernels", {}).get(wl_name)
    if e is None:
        return None, None, "no ncu capture of this workload in profiles/ncu_captures.json"
    sha = csrc_sha()
    if e.get("csrc_sha") != sha:
        return None, None, ("stale: the ncu capture %s was taken on kernel sources %s, these are %s" %
                            (e.get("file"), e.get("csrc_sha"), sha))
    return (e["dram_bytes_read"] + e["dram_bytes_write"], e.get("issue_active"),
            "ncu --set full, %s (kernel sources %s): dram__bytes_read.sum + dram__bytes_write.sum per launch" % (e.get("file"), sha))


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


def host_threads():
    """All the host threads this process may use.  Asked for explicitly: under torch.distributed.run the
    launcher exports OMP_NUM_THREADS=1, which must not turn the CPU arm into a one-thread run."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def strided_pixels(width, height, n_pixels, seed=1):
    """A bounded, uniformly spread sample of the frame's pixels (same workload mix as the frame)."""
    total = width * height
    n = min(n_pixels, total)
    rng = np.random.default_rng(seed)
    idx = np.sort(rng.choice(total, size=n, replace=False))
    return np.stack([idx % width, idx // width], axis=1).astype(np.int32)


def cpu_time_sample(ob, wl, target_seconds, as_written=True, probe_pixels=2048):
    """Time the oracle on a pixel sample sized for about target_seconds; returns (Mrays/s, description, cores, rays, s)."""
    kind, n, w, h, spp, depth = wl
    scene = make_scene(ob, kind, n, depth) if kind in ("builtin", "batch") else None
    if scene is None:
        import openglraytracer_b200 as rt   # scene GENERATOR only (host data); nothing is rendered by the product here
        scene = make_scene(rt, kind, n, depth)
    cam = ob.builtin_camera(0.0)
    cores = host_threads()
    probe = strided_pixels(w, h, probe_pixels)
    t0 = time.perf_counter()
    _, st = ob.render(scene, cam, w, h, spp=spp, pixels=probe, as_written=as_written, want_stats=True, threads=cores)
    dt = max(time.perf_counter() - t0, 1e-4)
    if dt < 0.6 * target_seconds:
        n_pix = int(min(w * h, max(probe_pixels, len(probe) * target_seconds / dt)))
        pix = strided_pixels(w, h, n_pix, seed=2)
        t0 = time.perf_counter()
        _, st = ob.render(scene, cam, w, h, spp=spp, pixels=pix, as_written=as_written, want_stats=True, threads=cores)
        dt = time.perf_counter() - t0
    else:
        pix = probe                         # the probe already is a sample of the wanted size
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
    per_step = 20.0 / max(args.steps + args.warmup, 1)
    per_step = min(max(per_step, 0.5), 6.0)
    for _ in range(args.warmup):
        cpu_time_sample(ob, wl, per_step)
    rays = 0
    secs = 0.0
    desc = ""
    cores = host_threads()
    for _ in range(args.steps):
        v, desc, cores, r, dt = cpu_time_sample(ob, wl, per_step)
        rays += r
        secs += dt
    value = rays / secs / 1e6
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / max(args.steps, 1),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": static_config(wl_name, wl, args.gpus),
        "note": "GL / llvmpipe unavailable on this box: oracle port of the shader, executed as written (per-test matrix "
                "rebuilds), %d host threads; the same CPU run whatever --gpus says" % cores,
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
    numa_mem = None
    if world > 1:
        from openglraytracer_b200.multigpu import bind_to_gpu_numa
        numa_cpus = bind_to_gpu_numa(local)         # before any page-locked allocation
        from openglraytracer_b200.multigpu import prefer_gpu_numa_memory
        numa_mem = prefer_gpu_numa_memory(local) if not os.environ.get("RT_NO_NUMA_MEMORY") else None
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

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def rank_max(x):
        if world == 1:
            return float(x)
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def time_kernel(cams_, w_, h_, spp_, out_ptr, steps, warmup, no_cull=0):
        """CUDA-event time of `steps` renders of the uploaded scene, one event pair per step, L2 flushed between."""
        for _ in range(warmup):
            tracer.render(None, cams_, w_, h_, spp=spp_, out_device_ptr=out_ptr, async_=True, stream=stream.cuda_stream,
                          no_cull=no_cull)
            flush.zero_()
        torch.cuda.synchronize(dev)
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        for a, b in evs:
            flush.zero_()                 # L2 flush between timed iterations (outside the event pair)
            a.record(stream)
            tracer.render(None, cams_, w_, h_, spp=spp_, out_device_ptr=out_ptr, async_=True, stream=stream.cuda_stream,
                          no_cull=no_cull)
            b.record(stream)
        torch.cuda.synchronize(dev)
        return [a.elapsed_time(b) for a, b in evs]

    # counters from the instrumented twin (untimed); the timed kernel is the plain one
    tracer.upload_scene(scene)
    _, st = tracer.render(None, cams, w, h, spp=spp, out_device_ptr=frame_dev.data_ptr(), want_stats=True)
    rays_per_frame = int(st["rays"])            # per launch of this rank (nfr frames in batch mode)
    rays_traversed = rays_per_frame - int(st["rays_not_traversed"])
    flops_per_frame = algorithmic_flops(st, nfr * w * h * spp)
    peak_tf, _ = rt.fp32_peak_tflops(local, 8192)

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    warm = max(args.warmup, 3)
    for _ in range(warm):
        tracer.render(None, cams, w, h, spp=spp, out_device_ptr=frame_dev.data_ptr(), async_=True, stream=stream.cuda_stream)
        flush.zero_()
    barrier()
    launches0 = rt.kernel_launch_count()
    t_region0 = time.perf_counter()
    step_ms = time_kernel(cams, w, h, spp, frame_dev.data_ptr(), args.steps, 0)
    barrier()
    t_region1 = time.perf_counter()
    launches = rt.kernel_launch_count() - launches0
    total_ms = rank_max(float(sum(step_ms)))
    ms_per_step = total_ms / args.steps
    rays_job = world * rays_per_frame           # frame mode: every rank renders its own copy of the frame (weak)
    trav_job = world * rays_traversed
    if batch and world > 1:                     # batch mode: the 256 poses are split over the ranks (strong)
        t = torch.tensor([rays_per_frame, rays_traversed], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        rays_job, trav_job = int(t[0].item()), int(t[1].item())
    value = rays_job / (ms_per_step * 1e-3) / 1e6
    value_trav = trav_job / (ms_per_step * 1e-3) / 1e6

    # e2e: the public call with host structs and a pinned host framebuffer
    pin = rt.PinnedImage((nfr, h, w, 4) if batch else (h, w, 4))
    scene_bytes = 132 * scene.objects_count + 60 * scene.scene.lights_count + 40 * nfr

    def time_e2e(scene_, cams_, w_, h_, spp_, out_arr, steps):
        for _ in range(3):
            tracer.render(scene_, cams_, w_, h_, spp=spp_, out=out_arr)
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            tracer.render(scene_, cams_, w_, h_, spp=spp_, out=out_arr)
        barrier()
        return rank_max(time.perf_counter() - t0)

    e2e_steps = max(3, min(args.steps, 50))
    e2e_s = time_e2e(scene, cams, w, h, spp, pin.array, e2e_steps)
    e2e_value = rays_job * e2e_steps / e2e_s / 1e6
    e2e_trav = trav_job * e2e_steps / e2e_s / 1e6
    checksum = int(pin.array[..., ::97, ::89, :].astype(np.uint64).sum())

    # ------------------------------------------------------------------ N = 1: BASELINE config 4 (scaled scene)
    scaled = None
    if world == 1 and not batch and not args.no_legs:
        scaled = {}
        ob = None if args.no_cpu_baseline else _oracle()
        for name in SCALED_LEGS:
            k2, n2, w2, h2, spp2, d2 = WORKLOADS[name]
            sc2 = make_scene(rt, k2, n2, d2)
            out2 = torch.empty((h2, w2, 4), dtype=torch.uint8, device=dev)
            pin2 = rt.PinnedImage((h2, w2, 4))
            tracer.upload_scene(sc2)
            leg = {"objects": n2, "width": w2, "height": h2, "spp": spp2, "max_depth": d2}
            ref_img = None
            shipped_only = name in SCALED_LEGS_SHIPPED_ONLY
            for label, mode in ((("shipped", 0),) if shipped_only else (("shipped", 0), ("every_slab_test_no_cull_1", 1))):
                _, s2 = tracer.render(None, cam, w2, h2, spp=spp2, out_device_ptr=out2.data_ptr(), want_stats=True, no_cull=mode)
                slow = mode == 1 and n2 > 4096
                ms = float(np.mean(time_kernel(cam, w2, h2, spp2, out2.data_ptr(), 3 if slow else 20, 1 if slow else 3, no_cull=mode)))
                fl = algorithmic_flops(s2, w2 * h2 * spp2)
                img = out2.cpu().numpy()
                if ref_img is None:
                    ref_img = img
                leg[label] = {"kernel_ms": ms, "value": s2["rays"] / (ms * 1e-3) / 1e6, "unit": UNIT,
                              "value_traversed": (s2["rays"] - s2["rays_not_traversed"]) / (ms * 1e-3) / 1e6,
                              "algorithmic_flops_per_launch": fl, "achieved_tflops": fl / (ms * 1e-3) / 1e12,
                              "frac_of_fp32_peak": fl / (ms * 1e-3) / 1e12 / peak_tf,
                              "box_tests": int(s2["box_tests"]), "sphere_tests": int(s2["sphere_tests"]),
                              "bounding_sphere_tests": int(s2["cull_tests"]), "packet_cone_tests": int(s2["packet_tests"]),
                              "byte_identical_to_shipped": bool(np.array_equal(img, ref_img))}
            sc2p = sc2.pinned()           # inputs in page-locked host memory (the e2e contract), re-uploaded every step
            e2s = time_e2e(sc2p, cam, w2, h2, spp2, pin2.array, 10)
            leg["e2e"] = {"ms_per_step": 1e3 * e2s / 10, "value": s2["rays"] * 10 / e2s / 1e6, "unit": UNIT,
                          "h2d_bytes_per_step": 132 * n2 + 180 + 40, "d2h_bytes_per_step": w2 * h2 * 4,
                          "frame_equals_device_frame": bool(np.array_equal(pin2.array, ref_img))}
            e2p = time_e2e(sc2, cam, w2, h2, spp2, pin2.array, 5)
            leg["e2e"]["ms_per_step_pageable_objects"] = 1e3 * e2p / 5
            tracer.upload_scene(sc2)
            sc2p.free_pinned()
            if ob is not None and not shipped_only:
                v, desc, cores, _, _ = cpu_time_sample(ob, WORKLOADS[name], 4.0, probe_pixels=512 if n2 > 4096 else 2048)
                leg["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc}
            pin2.free()
            del out2
            scaled[name] = leg
        tracer.upload_scene(scene)
        scaled["note"] = ("BASELINE config 4 at 1920x1080, 1 spp, depth 0 (and the shipped kernels at depth 2), kernel time with "
                          "inputs resident (CUDA events, L2 flushed between steps).  every_slab_test_no_cull_1 is the brute-force loop with all 59-flop "
                          "slab tests executed -- the loop north_star's >= 60 %% of FP32 peak is stated on; shipped puts "
                          "conservative bounding / packet tests in front of it (same bytes of image, flag above) and its "
                          "fraction is a fraction of the far smaller flop count it executes.  FP32 peak %.1f TFLOP/s measured "
                          "in this run." % peak_tf)

    # ------------------------------------------------------------------ N > 1: one frame split into contiguous row bands
    def band_modes(w_, h_, spp_, cam_, single_frame, n_iter):
        """NCCL gather and fused peer-store gather of one frame's row bands; every rank calls this."""
        out = {}
        cost_bands = balanced_row_bands(probe_row_cost(tracer, None, cam_, w_, h_), world)
        torch.cuda.synchronize(dev)
        br = BandRenderer(tracer, w_, h_, rank, world, dev, bands=cost_bands)
        for _ in range(3):
            br.render_and_gather(None, cam_, spp=spp_)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(n_iter):
            full = br.render_and_gather(None, cam_, spp=spp_)
        e1.record(stream)
        barrier()
        ms_b = rank_max(e0.elapsed_time(e1)) / n_iter
        same = bool(np.array_equal(full.cpu().numpy(), single_frame)) if rank == 0 else True
        out["nccl_gather"] = {"ms_per_frame": ms_b, "collective": "one grouped ncclSend/ncclRecv exchange of the exact band "
                              "sizes, received in place in rank 0's frame", "bands": str(cost_bands),
                              "byte_identical_to_single_gpu": same}
        del br
        try:
            for label, bset in (("peer_store", cost_bands), ("peer_store_feedback_bands", None)):
                pf = PeerFrame(tracer, w_, h_, rank, world, dev, bands=bset or cost_bands)
                if bset is None:     # three feedback steps on the measured per-band times
                    bset = pf.autotune(None, cam_, cost_bands, iters=3, spp=spp_)
                band_ms = pf.measure_bands(None, cam_, spp=spp_)
                for _ in range(3):
                    pf.render(None, cam_, spp=spp_)
                barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                for _ in range(n_iter):
                    full = pf.render(None, cam_, spp=spp_)
                e1.record(stream)
                barrier()
                # the frame is complete on rank 0's stream when its wait kernel has passed: rank 0's clock is the frame's
                ms_p = rank_max(e0.elapsed_time(e1)) / n_iter
                same_p = bool(np.array_equal(full.cpu().numpy(), single_frame)) if rank == 0 else True
                timed_out = rt.api.wait_flags_timed_out()
                pf.close()
                out[label] = {"ms_per_frame": ms_p,
                              "collective": "none: kernels store into rank 0's frame over NVLink (CUDA IPC mapping); a one-thread "
                                            "kernel behind each rank's render raises a completion word there, one warp on rank 0 waits",
                              "bands": str(bset), "band_kernel_ms": band_ms,
                              "imbalance_max_over_mean": max(band_ms) / (sum(band_ms) / len(band_ms)),
                              "completion_and_launch_ms": ms_p - max(band_ms),
                              "byte_identical_to_single_gpu": same_p, "wait_timed_out": bool(timed_out)}
        except Exception as exc:   # IPC refused (container policy): the NCCL gather above stands
            out["peer_store"] = {"unavailable": repr(exc)[:200]}
        return out

    bands = None
    bands4k = None
    batch256 = None
    if world > 1 and not batch and not args.no_legs:
        n_b = max(3, min(args.steps, 50))
        bands = band_modes(w, h, spp, cam, pin.array, n_b)
        bands["rays_per_frame"] = rays_per_frame
        bands["single_gpu_ms"] = ms_per_step
        for k in ("nccl_gather", "peer_store", "peer_store_feedback_bands"):
            if k in bands and "ms_per_frame" in bands[k]:
                bands[k]["value"] = rays_per_frame / (bands[k]["ms_per_frame"] * 1e-3) / 1e6
                bands[k]["speedup_vs_single_gpu"] = ms_per_step / bands[k]["ms_per_frame"]
        bands.update({"unit": UNIT, "scaling": "strong"})

    if world > 1 and wl_name == DEFAULT_WORKLOAD and not args.no_legs:
        # BASELINE config 3: 3840x2160, 16 jittered spp, depth 4, row bands over the ranks
        _, _, w3, h3, spp3, _ = WORKLOADS["builtin_4k_spp16_d4"]
        single = np.zeros((h3, w3, 4), np.uint8)
        out3 = torch.empty((h3, w3, 4), dtype=torch.uint8, device=dev)
        _, s3 = tracer.render(None, cam, w3, h3, spp=spp3, out_device_ptr=out3.data_ptr(), want_stats=True)
        ms3 = rank_max(float(np.mean(time_kernel(cam, w3, h3, spp3, out3.data_ptr(), 3, 1))))
        if rank == 0:
            single = out3.cpu().numpy()
        del out3
        bands4k = band_modes(w3, h3, spp3, cam, single, 5)
        bands4k.update({"width": w3, "height": h3, "spp": spp3, "max_depth": depth, "rays_per_frame": int(s3["rays"]),
                        "single_gpu_ms": ms3, "unit": UNIT, "scaling": "strong"})
        for k in ("nccl_gather", "peer_store", "peer_store_feedback_bands"):
            if k in bands4k and "ms_per_frame" in bands4k[k]:
                bands4k[k]["value"] = int(s3["rays"]) / (bands4k[k]["ms_per_frame"] * 1e-3) / 1e6
                bands4k[k]["speedup_vs_single_gpu"] = ms3 / bands4k[k]["ms_per_frame"]

        # BASELINE config 5: 256 turntable poses at 1024 x 1024, whole frames per rank (k = rank mod world)
        _, _, w5, h5, spp5, d5 = WORKLOADS["batch256_1024_d4"]
        mine = frame_stride_set(rank, world, BATCH_FRAMES)
        poses = rt.turntable_cameras(BATCH_FRAMES)
        cams5 = (rt.Camera * len(mine))(*[poses[k] for k in mine])
        out5 = torch.empty((len(mine), h5, w5, 4), dtype=torch.uint8, device=dev)
        _, s5 = tracer.render(None, cams5, w5, h5, spp=spp5, out_device_ptr=out5.data_ptr(), want_stats=True)
        barrier()
        ms5 = rank_max(float(np.mean(time_kernel(cams5, w5, h5, spp5, out5.data_ptr(), 3, 1))))
        t = torch.tensor([float(s5["rays"])], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        pin5 = rt.PinnedImage((len(mine), h5, w5, 4))
        e5 = time_e2e(scene, cams5, w5, h5, spp5, pin5.array, 3)
        same5 = bool(np.array_equal(pin5.array[0], tracer.render(None, poses[mine[0]], w5, h5, spp=spp5)))
        pin5.free()
        del out5
        batch256 = {"frames": BATCH_FRAMES, "width": w5, "height": h5, "spp": spp5, "max_depth": d5,
                    "split": "frame k on rank k mod %d, no exchange" % world, "ms_per_batch": ms5,
                    "value": float(t.item()) / (ms5 * 1e-3) / 1e6, "unit": UNIT, "scaling": "strong",
                    "frames_per_s": BATCH_FRAMES / (ms5 * 1e-3),
                    "e2e": {"ms_per_batch": 1e3 * e5 / 3, "value": float(t.item()) * 3 / e5 / 1e6,
                            "d2h_bytes_per_step_per_rank": len(mine) * w5 * h5 * 4},
                    "first_frame_equals_single_call": same5}

    clocks = sampler.stop(t_region0, t_region1) if rank == 0 else None
    if rank == 0:
        kernel_ms = float(np.mean(step_ms))
        achieved = flops_per_frame / (kernel_ms * 1e-3) / 1e12
        traffic, issue, traffic_note = ncu_capture(wl_name)
        cfg = static_config(wl_name, wl, world)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": warm, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong" if batch else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": cfg,
            "value_traversed": value_trav,
            "rays": {"per_frame_reference_equivalent": rays_per_frame // nfr, "per_frame_traversed": rays_traversed // nfr,
                     "note": "value counts the rays the reference traces (one per get_closest_collision call, glsl:738); "
                             "value_traversed only those the kernel traverses: shadow rays toward a light without diffuse / "
                             "specular terms, or from a surface whose Phong weight is 0, cannot change the pixel and are skipped"},
            "kernel": "fast (fmad on, no fast-math); kernel sources " + csrc_sha(),
            "host_binding": (("each rank on the %d cores NVML names for its GPU" % len(numa_cpus)) if numa_cpus else "launcher default") +
                            ("; page-locked frames preferred on the GPU's NUMA node (rank 0: node %d)" % numa_mem if numa_mem is not None
                             else "; memory policy unchanged (one node, unknown topology or set_mempolicy refused)"),
            "e2e": {"value": e2e_value, "value_traversed": e2e_trav, "unit": UNIT, "h2d_bytes_per_step": scene_bytes,
                    "d2h_bytes_per_step": nfr * w * h * 4, "steps": e2e_steps, "ms_per_step": 1e3 * e2e_s / e2e_steps,
                    "checksum": checksum},
            "gpu_launches": int(launches),
            "roofline": {"bound": "fp32", "achieved": achieved, "peak": peak_tf, "unit": "TFLOP/s",
                         "frac": achieved / peak_tf, "traffic": traffic, "issue_slots_active_ncu": issue,
                         "traffic_note": traffic_note + "; algorithmic bytes = the %d-byte frame + the scene records: "
                                         "FP32-issue bound, not HBM" % (nfr * w * h * 4),
                         "peak_source": "FP32 FMA-chain microbenchmark in this run (MEASURED_PEAKS.json has no FP32 "
                                        "figure); nominal 148 SM x 128 x 2 x 1.965 GHz = 74.4",
                         "algorithmic_flops_per_launch": flops_per_frame, "kernel_ms": kernel_ms,
                         "box_tests": int(st["box_tests"]), "sphere_tests": int(st["sphere_tests"]),
                         "bounding_sphere_tests": int(st["cull_tests"]),
                         "packet_cone_tests": int(st.get("packet_tests", 0)),
                         "hit_nodes": int(st["hit_nodes"]), "shaded_lights": int(st["shaded_lights"]),
                         "shadow_rays_traversed": int(st["shadow_rays"] - st["rays_not_traversed"]),
                         "lane_use": st["nodes"] / max(1, 32 * st["warp_iters"]),
                         "flops_per_unit": FLOPS_NOTE},
            "clocks": clocks,
        }
        if scaled:
            line["scaled"] = scaled
        if bands:
            line["bands_gather"] = bands
        if bands4k:
            line["bands_gather_4k_spp16"] = bands4k
        if batch256:
            line["batch256"] = batch256
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
    ap.add_argument("--no-legs", action="store_true", help="skip the extra legs (scaled / bands_gather* / batch256)")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, args.workload, wl)
    else:
        run_gpu(args, args.workload, wl)


if __name__ == "__main__":
    main()
