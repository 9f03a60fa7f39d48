"""Parity metric of north_star: <= 1 LSB per RGBA8 channel on >= 99.9 % of pixels;
larger diffs are counted and located."""
import numpy as np

TOL_LSB = 1            # per channel
MIN_FRACTION = 0.999   # of pixels within TOL_LSB


def compare(gpu, ref):
    gpu = np.asarray(gpu)
    ref = np.asarray(ref)
    assert gpu.shape == ref.shape, (gpu.shape, ref.shape)
    d = np.abs(gpu.astype(np.int16) - ref.astype(np.int16)).reshape(-1, gpu.shape[-1])
    per_pixel = d.max(axis=1)
    n = per_pixel.size
    rep = {
        "pixels": int(n),
        "exact": int((per_pixel == 0).sum()),
        "within_1": int((per_pixel <= TOL_LSB).sum()),
        "over_1": int((per_pixel > TOL_LSB).sum()),
        "max_diff": int(per_pixel.max()) if n else 0,
        "frac_within_1": float((per_pixel <= TOL_LSB).mean()) if n else 1.0,
        "frac_exact": float((per_pixel == 0).mean()) if n else 1.0,
    }
    return rep


def assert_parity(gpu, ref, what=""):
    rep = compare(gpu, ref)
    assert rep["frac_within_1"] >= MIN_FRACTION, "%s parity failed: %r" % (what, rep)
    return rep


def edge_mask(ref, radius=1):
    """Pixels whose neighbourhood has a colour discontinuity in the oracle image:
    silhouettes / shadow edges, where an FP hit decision may legitimately flip."""
    r = ref[..., :3].astype(np.int16)
    h, w = r.shape[:2]
    m = np.zeros((h, w), bool)
    for dy in range(-radius, radius + 1):
        for dx in range(-radius, radius + 1):
            if dx == 0 and dy == 0:
                continue
            sh = np.roll(np.roll(r, dy, axis=0), dx, axis=1)
            m |= (np.abs(sh - r).max(axis=2) > 8)
    return m
