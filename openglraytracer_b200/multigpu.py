"""Multi-GPU sharding of the ray-scene kernel: one process per GPU.

Pixels are independent (the only write is imageStore at glsl:404), so a frame
shards as contiguous row bands with NO data-path collective; the one exchange
is the gather of the finished bands on rank 0 (north_star).  Batch mode shards
whole frames and needs no exchange at all.

torch is plumbing only: device buffers, streams and torch.distributed (NCCL
over NVLink on the GPU box, gloo in the CPU tests).
"""
import os

import torch
import torch.distributed as dist


def bind_to_gpu_numa(device_index):
    """One process per GPU: run this process on the CPU cores next to its GPU (NVML's ideal CPU affinity), so
    that the page-locked host frame the kernel stores into is allocated on the GPU's own NUMA node and the
    posted PCIe writes do not cross sockets.  Returns the cores bound to, or None when NVML is not usable (the
    process then stays where the launcher put it)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        n_cpu = os.cpu_count() or 1
        handle = pynvml.nvmlDeviceGetHandleByIndex(int(device_index))
        words = pynvml.nvmlDeviceGetCpuAffinity(handle, (n_cpu + 63) // 64)
        cpus = {64 * i + b for i, wd in enumerate(words) for b in range(64) if (int(wd) >> b) & 1}
        cpus &= set(os.sched_getaffinity(0))
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return sorted(cpus)
    except Exception:        # noqa: BLE001 - best effort: no NVML, no permission, unknown topology
        return None


def gpu_numa_node(device_index):
    """NUMA node the GPU's PCIe slot hangs off (sysfs), or None."""
    try:
        props = torch.cuda.get_device_properties(int(device_index))
        bus = "%04x:%02x:%02x.0" % (props.pci_domain_id, props.pci_bus_id, props.pci_device_id)
        with open("/sys/bus/pci/devices/%s/numa_node" % bus) as f:
            node = int(f.read().strip())
        return node if node >= 0 else None
    except Exception:        # noqa: BLE001 - attribute / file missing: unknown topology
        return None


def prefer_gpu_numa_memory(device_index):
    """Page-locked frames of this process on the GPU's OWN NUMA node: set_mempolicy(MPOL_PREFERRED, node) before the
    frames are allocated.  The CPU affinity NVML reports can be one mask for every GPU (a container's cpuset on one
    socket): the processes then all allocate on that socket, and the posted PCIe frame stores of the GPUs behind the
    other socket cross the socket interconnect.  Memory policy is independent of where the threads run.  Best effort:
    returns the node, or None (unknown topology, single node, or the sandbox refuses the system call)."""
    import ctypes
    node = gpu_numa_node(device_index)
    if node is None or not os.path.isdir("/sys/devices/system/node/node%d" % node):
        return None
    try:
        n_nodes = len([d for d in os.listdir("/sys/devices/system/node") if d.startswith("node") and d[4:].isdigit()])
        if n_nodes < 2:
            return None
        libc = ctypes.CDLL(None, use_errno=True)
        mask = (ctypes.c_ulong * 16)()
        mask[node // 64] = 1 << (node % 64)
        MPOL_PREFERRED, SYS_set_mempolicy = 1, 238          # x86_64
        if os.uname().machine != "x86_64":
            return None
        rc = libc.syscall(SYS_set_mempolicy, MPOL_PREFERRED, mask, ctypes.c_ulong(16 * 64))
        return node if rc == 0 else None
    except Exception:        # noqa: BLE001
        return None


def row_band(rank, world, height):
    """Contiguous band [y0, y1) of rank `rank`: rows g*H/G .. (g+1)*H/G."""
    return (rank * height) // world, ((rank + 1) * height) // world


def frame_block(rank, world, n_frames):
    """Contiguous block of frames [k0, k1) of rank `rank` (batch mode)."""
    return (rank * n_frames) // world, ((rank + 1) * n_frames) // world


def balanced_row_bands(row_cost, world, align=4):
    """Contiguous bands of (nearly) equal COST: row_cost[y] is any per-row cost estimate (rays,
    time ...).  Boundaries are multiples of `align` rows (the kernel's tiles are 4 rows high), every
    band keeps at least one row, and the bands partition [0, H).  Deterministic: every rank that
    feeds the same costs gets the same bands, so no exchange is needed.  Floor and glass rows cost
    more rays than wall rows (SURVEY §7 hard part 5); equal-height bands leave GPUs idle."""
    h = len(row_cost)
    if world >= h:
        return [row_band(r, world, h) for r in range(world)]
    total = float(sum(row_cost))
    if total <= 0.0:
        return [row_band(r, world, h) for r in range(world)]
    cuts = [0]
    acc, y = 0.0, 0
    for r in range(1, world):
        target = total * r / world
        while y < h and acc + row_cost[y] * 0.5 < target:
            acc += row_cost[y]
            y += 1
        c = int(round(y / align)) * align
        c = max(c, cuts[-1] + 1)
        c = min(c, h - (world - r))
        cuts.append(c)
    cuts.append(h)
    return [(cuts[r], cuts[r + 1]) for r in range(world)]


def refine_row_bands(bands, band_ms, height, align=4):
    """One feedback step: given the measured render time of every band, re-cut the frame so that the
    bands would take equal time if cost were uniform inside each old band (piecewise-constant cost
    density band_ms[r] / rows_r).  Two or three steps on a static view converge; in an animation the
    previous frame's times play the same role."""
    cost = [0.0] * height
    for (a, b), ms in zip(bands, band_ms):
        for y in range(a, b):
            cost[y] = float(ms) / max(1, b - a)
    return balanced_row_bands(cost, len(bands), align)


def probe_row_cost(raytracer, scene, camera, width, height, spp=1, strips=72, shrink=8):
    """Per-row cost estimate for balanced_row_bands: the reference ray count (rt_stats.rays) of
    `strips` horizontal strips rendered at 1/shrink of the width with the counting kernel (a few
    tiny launches; the counts are deterministic, so every rank computes the same numbers)."""
    import numpy as np
    strips = max(1, min(strips, height))
    w_probe = max(8, width // shrink)
    cost = np.zeros(height, np.float64)
    if scene is not None:
        raytracer.upload_scene(scene)
    for s in range(strips):
        y0, y1 = (s * height) // strips, ((s + 1) * height) // strips
        # one row in the middle of the strip stands for the strip; absolute pixel rows keep the geometry
        ym = (y0 + y1) // 2
        _, st = raytracer.render(None, camera, w_probe, height, spp=1, y0=ym, y1=ym + 1, band_only=True,
                                 want_stats=True)
        cost[y0:y1] = st["rays"] + 0.25 * st["cull_tests"] / max(1, w_probe)
    return cost


def frame_stride_set(rank, world, n_frames):
    """Frames k = rank (mod world) (batch mode): neighbouring turntable poses cost about the same, so
    the interleaved split is balanced where contiguous blocks are not (7.45x -> see profiles/ at 8 GPUs)."""
    return list(range(rank, n_frames, world))


def gather_bands(band, height, width, group=None, dst=0, bands=None, out=None):
    """Gather every rank's band (uint8 [rows_r, W, 4]) into the full frame on `dst`.

    One grouped exchange of point-to-point operations of the EXACT band sizes: every other rank sends its
    band, `dst` receives each one straight into its rows of the frame (`out`: a persistent uint8 [H, W, 4]
    tensor, allocated when missing).  No padding to the tallest band, no staging copies after the
    collective (NCCL 2.27 has no ncclGather; a dist.gather of padded bands plus eight slice copies was
    slower at 8 GPUs than at 4).  Returns the frame on dst, None elsewhere.
    """
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    if bands is None:
        bands = [row_band(r, world, height) for r in range(world)]
    y0, y1 = bands[rank]
    assert band.shape == (y1 - y0, width, 4) and band.dtype == torch.uint8, (band.shape, band.dtype)

    def peer(r):        # P2POp wants global ranks
        return dist.get_global_rank(group, r) if group is not None else r

    if rank == dst:
        frame = out if out is not None else torch.empty((height, width, 4), dtype=torch.uint8, device=band.device)
        mine = frame[y0:y1]
        if mine.data_ptr() != band.data_ptr():
            mine.copy_(band)
        ops = [dist.P2POp(dist.irecv, frame[bands[r][0]:bands[r][1]], peer(r), group)
               for r in range(world) if r != dst and bands[r][1] > bands[r][0]]
    else:
        frame = None
        ops = [dist.P2POp(dist.isend, band.contiguous(), peer(dst), group)] if y1 > y0 else []
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    return frame


class BandRenderer:
    """Row-band renderer of one rank: renders rows [y0, y1) of every frame into a
    device tensor through the C ABI (device output, caller's stream).  On rank `dst` the band IS the
    rank's slice of a persistent full frame, so its own rows need no copy."""

    def __init__(self, raytracer, width, height, rank, world, device, bands=None, dst=0):
        self.rt = raytracer
        self.width, self.height = width, height
        self.rank, self.world, self.dst = rank, world, dst
        self.bands = bands if bands is not None else [row_band(r, world, height) for r in range(world)]
        self.y0, self.y1 = self.bands[rank]
        self.frame = None
        if rank == dst:
            self.frame = torch.empty((height, width, 4), dtype=torch.uint8, device=device)
            self.band = self.frame[self.y0:self.y1]
        else:
            self.band = torch.empty((self.y1 - self.y0, width, 4), dtype=torch.uint8, device=device)
        # A real stream handle: the C ABI reads a NULL stream as "the context's own stream",
        # and torch's legacy default stream IS the NULL handle.
        self.stream = torch.cuda.Stream(device)

    def render(self, scene, camera, **kw):
        """Launch the band render on this renderer's stream (asynchronous), ordered after
        whatever the caller's current stream has queued."""
        self.stream.wait_stream(torch.cuda.current_stream(self.band.device))
        stream = self.stream.cuda_stream
        self.rt.render(scene, camera, self.width, self.height, y0=self.y0, y1=self.y1, band_only=True,
                       out_device_ptr=self.band.data_ptr(), async_=True, stream=stream, **kw)
        return self.band

    def render_and_gather(self, scene, camera, group=None, **kw):
        band = self.render(scene, camera, **kw)
        with torch.cuda.stream(self.stream):
            frame = gather_bands(band, self.height, self.width, group=group, dst=self.dst, bands=self.bands,
                                 out=self.frame)
        torch.cuda.current_stream(self.band.device).wait_stream(self.stream)
        return frame


class _RawCudaBuffer:
    """A cudaMalloc'ed buffer exposed through __cuda_array_interface__ so that torch can view it."""

    def __init__(self, ptr, shape, typestr="|u1"):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2, "strides": None}


class PeerFrame:
    """Fused gather (SURVEY §5 / §8e stage 2): rank `dst` owns the frame; every other rank maps
    it through a CUDA IPC handle and its render kernel's 16-byte framebuffer stores
    (rt_kernel.cuh, glsl:404 imageStore) land directly in that buffer over NVLink.

    No collective on the frame's path at all.  Behind the frame sit world + 1 32-bit words in the same
    allocation: word r is raised to the frame's epoch by a one-thread kernel queued behind rank r's render kernel
    (rt_options.done_flag: every store of the render is visible at the kernel boundary, the word follows with
    system-scope release), and `dst` puts one warp on its stream that waits for all of them (rt_wait_flags).  The last word runs the other way: `dst` raises it to
    k - 1 when the caller asks for frame k (the previous frame is then his to overwrite), and every producer's
    stream waits for it before its kernel may store -- so a fast rank never writes into a frame `dst` still
    reads.  Setup (IPC handle exchange) uses torch.distributed once; frames do not."""

    def __init__(self, raytracer, width, height, rank, world, device, group=None, dst=0, bands=None):
        import ctypes as C
        from . import _abi
        self.rt, self.width, self.height = raytracer, width, height
        self.rank, self.world, self.dst, self.group = rank, world, dst, group
        self.y0, self.y1 = (bands if bands is not None else [row_band(r, world, height) for r in range(world)])[rank]
        self.device = device
        self.stream = torch.cuda.Stream(device)
        self._lib = _abi.lib()
        nbytes = width * height * 4
        self._flags_off = (nbytes + 255) // 256 * 256
        total = self._flags_off + 4 * (world + 1) + 256
        handle = torch.zeros(64, dtype=torch.uint8, device=device)
        self._own = C.c_void_p()
        self._mapped = C.c_void_p()
        self.epoch = 0
        # Every rank goes through the same collectives whatever fails locally; the outcome is agreed
        # with an all-reduce so that either all ranks own a working PeerFrame or all raise.
        err = None
        self.ptr = None
        if rank == dst:
            try:
                _abi.check(self._lib.rt_device_alloc(C.byref(self._own), total), "rt_device_alloc")
                buf = C.create_string_buffer(64)
                _abi.check(self._lib.rt_ipc_export(self._own, buf), "rt_ipc_export")
                handle.copy_(torch.frombuffer(bytearray(buf.raw), dtype=torch.uint8))
                self.ptr = self._own.value
                torch.as_tensor(_RawCudaBuffer(self.ptr + self._flags_off, (4 * (world + 1),)), device=device).zero_()
                torch.cuda.synchronize(device)
            except Exception as exc:
                err = exc
        dist.broadcast(handle, src=dst, group=group)
        if rank != dst:
            try:
                raw = bytes(handle.cpu().numpy().tobytes())
                _abi.check(self._lib.rt_ipc_import(raw, C.byref(self._mapped)), "rt_ipc_import")
                self.ptr = self._mapped.value
            except Exception as exc:
                err = exc
        ok = torch.tensor([0 if err else 1], dtype=torch.int32, device=device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
        if int(ok.item()) == 0:
            if self._mapped:
                self._lib.rt_ipc_close(self._mapped)
            if self._own:
                self._lib.rt_device_free(self._own)
            raise RuntimeError("PeerFrame: CUDA IPC mapping of rank %d's frame failed on some rank (%r)" % (dst, err))
        self.frame = None
        if rank == dst:
            self.frame = torch.as_tensor(_RawCudaBuffer(self.ptr, (height, width, 4)), device=device)
        self._flags = self.ptr + self._flags_off              # word r: rank r's band of the current epoch is complete
        self._consumed = self._flags + 4 * world              # raised by dst: frames up to this epoch may be overwritten

    def render(self, scene, camera, **kw):
        """All ranks: render this rank's band of the next frame into dst's frame.  Returns the full frame (uint8
        [H, W, 4]) on dst -- valid for work queued on the caller's current stream, until the next render() call,
        which hands it back to the producers -- and None elsewhere.  No host synchronisation, no collective."""
        from . import api
        cur = torch.cuda.current_stream(self.device)
        self.stream.wait_stream(cur)
        self.epoch += 1
        st = self.stream.cuda_stream
        if self.rank == self.dst:
            # whatever the caller queued on `cur` that reads the previous frame is ordered before this
            _check(self._lib.rt_write_flag(self._consumed, (self.epoch - 1) & 0xFFFFFFFF, st), "rt_write_flag")
        else:
            api.wait_flags(self._consumed, 1, self.epoch - 1, stream=st)
        self.rt.render(scene, camera, self.width, self.height, y0=self.y0, y1=self.y1, band_only=False,
                       out_device_ptr=self.ptr, async_=True, stream=st,
                       done_flag=self._flags + 4 * self.rank, done_value=self.epoch, **kw)
        if self.rank == self.dst:
            api.wait_flags(self._flags, self.world, self.epoch, stream=st)
        cur.wait_stream(self.stream)
        return self.frame

    def set_bands(self, bands):
        self.y0, self.y1 = bands[self.rank]

    def autotune(self, scene, camera, bands, iters=3, **kw):
        """Feedback balancing of the contiguous bands: render, all-gather the per-rank band times
        (CUDA events on the render stream), re-cut (refine_row_bands); returns the final bands.  The
        time exchange is a tiny all-gather outside the frame's data path."""
        bands = list(bands)
        best = None                     # (slowest band's ms, bands): band times of a 0.1 ms frame are noisy, keep the best cut seen
        for it in range(iters + 1):
            self.set_bands(bands)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            cur = torch.cuda.current_stream(self.device)
            self.stream.wait_stream(cur)
            for k in range(3):          # first pass warms up, two are timed
                if k == 1:
                    e0.record(self.stream)
                self.rt.render(scene, camera, self.width, self.height, y0=self.y0, y1=self.y1, band_only=False,
                               out_device_ptr=self.ptr, async_=True, stream=self.stream.cuda_stream, **kw)
            e1.record(self.stream)
            self.stream.synchronize()
            mine = torch.tensor([e0.elapsed_time(e1) / 2], dtype=torch.float32, device=self.device)
            every = [torch.zeros_like(mine) for _ in range(self.world)]
            dist.all_gather(every, mine, group=self.group)
            self.band_ms = [float(t.item()) for t in every]
            if best is None or max(self.band_ms) < best[0]:
                best = (max(self.band_ms), list(bands))
            if it < iters:
                bands = refine_row_bands(bands, self.band_ms, self.height)
        bands = best[1]
        self.set_bands(bands)
        return bands

    def measure_bands(self, scene, camera, **kw):
        """Per-band kernel time (ms) of the current bands, one entry per rank (a tiny all-gather)."""
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.stream.wait_stream(torch.cuda.current_stream(self.device))
        for k in range(2):
            if k == 1:
                e0.record(self.stream)
            self.rt.render(scene, camera, self.width, self.height, y0=self.y0, y1=self.y1, band_only=False,
                           out_device_ptr=self.ptr, async_=True, stream=self.stream.cuda_stream, **kw)
        e1.record(self.stream)
        self.stream.synchronize()
        mine = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float32, device=self.device)
        every = [torch.zeros_like(mine) for _ in range(self.world)]
        dist.all_gather(every, mine, group=self.group)
        return [float(t.item()) for t in every]

    def close(self):
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)
        if self._mapped:
            self._lib.rt_ipc_close(self._mapped)
            self._mapped = None
        dist.barrier(group=self.group)
        if self._own:
            self.frame = None
            self._lib.rt_device_free(self._own)
            self._own = None


def _check(code, where):
    from . import _abi
    _abi.check(code, where)
