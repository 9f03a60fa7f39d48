"""Multi-GPU sharding of the ray-scene kernel: one process per GPU.

Pixels are independent (the only write is imageStore at glsl:404), so a frame
shards as contiguous row bands with NO data-path collective; the one exchange
is the gather of the finished bands on rank 0 (north_star).  Batch mode shards
whole frames and needs no exchange at all.

torch is plumbing only: device buffers, streams and torch.distributed (NCCL
over NVLink on the GPU box, gloo in the CPU tests).
"""
import torch
import torch.distributed as dist


def row_band(rank, world, height):
    """Contiguous band [y0, y1) of rank `rank`: rows g*H/G .. (g+1)*H/G."""
    return (rank * height) // world, ((rank + 1) * height) // world


def frame_block(rank, world, n_frames):
    """Contiguous block of frames [k0, k1) of rank `rank` (batch mode)."""
    return (rank * n_frames) // world, ((rank + 1) * n_frames) // world


def gather_bands(band, height, width, group=None, dst=0):
    """Gather every rank's band (uint8 [rows_r, W, 4]) into the full frame on `dst`.

    Bands may differ by one row when H % G != 0; they are padded to the largest
    band for the collective and trimmed on arrival.  Returns uint8 [H, W, 4] on
    dst, None elsewhere.
    """
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    max_rows = max(row_band(r, world, height)[1] - row_band(r, world, height)[0] for r in range(world))
    y0, y1 = row_band(rank, world, height)
    assert band.shape == (y1 - y0, width, 4) and band.dtype == torch.uint8, (band.shape, band.dtype)
    if y1 - y0 == max_rows:
        send = band.contiguous()
    else:
        send = torch.zeros((max_rows, width, 4), dtype=torch.uint8, device=band.device)
        send[: y1 - y0] = band
    if rank == dst:
        parts = [torch.empty_like(send) for _ in range(world)]
        dist.gather(send, parts, dst=dst, group=group)
        frame = torch.empty((height, width, 4), dtype=torch.uint8, device=band.device)
        for r in range(world):
            a, b = row_band(r, world, height)
            frame[a:b] = parts[r][: b - a]
        return frame
    dist.gather(send, None, dst=dst, group=group)
    return None


class BandRenderer:
    """Row-band renderer of one rank: renders rows [y0, y1) of every frame into a
    device tensor through the C ABI (device output, caller's stream)."""

    def __init__(self, raytracer, width, height, rank, world, device):
        self.rt = raytracer
        self.width, self.height = width, height
        self.rank, self.world = rank, world
        self.y0, self.y1 = row_band(rank, world, height)
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
            frame = gather_bands(band, self.height, self.width, group=group)
        torch.cuda.current_stream(self.band.device).wait_stream(self.stream)
        return frame
