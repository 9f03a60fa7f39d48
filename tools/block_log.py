"""Per-block timing of the CTA-level packet kernel (render_block_kernel).

RT_BLOCK_LOG=path makes the instrumented twin write one 16-word record per block:
  [0,1] start (globaltimer ns)  [2] ns until list A is complete  [3] list A entries  [4] list B entries
  [5] flags: 1 list B used (always; 2 = not, older logs), 4 use_a, 8 !use_a, bits 4-6 light cones on, bits 8-10 balls
  [6] SM id  [7] CTA  [8..15] ns from block start until warp k reached the end of its work item
Usage: python tools/block_log.py n_objects [width height depth]   |   python tools/block_log.py --analyze log.bin
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    if len(sys.argv) > 2 and sys.argv[1] == "--analyze":          # summarise an existing log (no GPU needed)
        return analyze(sys.argv[2])
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
    w = int(sys.argv[2]) if len(sys.argv) > 2 else 1920
    h = int(sys.argv[3]) if len(sys.argv) > 3 else 1080
    depth = int(sys.argv[4]) if len(sys.argv) > 4 else 0
    import openglraytracer_b200 as rt
    path = os.path.join(ROOT, "gpurun_out", "block_log_%d.bin" % n)
    os.makedirs(os.path.dirname(path), exist_ok=True)
    os.environ["RT_BLOCK_LOG"] = path
    sc = rt.scaled_scene(n, max_depth=depth)
    cam = rt.builtin_camera(0.0)
    with rt.Raytracer(0) as r:
        r.render(sc, cam, w, h)
        img, st = r.render(sc, cam, w, h, want_stats=True)
    print("kernel_ms", st["kernel_ms"])
    analyze(path)


def analyze(path):
    a = np.fromfile(path, dtype=np.uint32).reshape(-1, 16)
    t0 = a[:, 0].astype(np.uint64) | (a[:, 1].astype(np.uint64) << np.uint64(32))
    ok = t0 > 0
    a, t0 = a[ok], t0[ok]
    warp = a[:, 8:16].astype(np.float64) / 1e3            # us
    dur = warp.max(axis=1)
    start = (t0 - t0.min()).astype(np.float64) / 1e3
    end = start + dur
    print("blocks", len(a), "twin span us", end.max(), "sum of block us", dur.sum(), "per CTA mean", dur.sum() / len(set(a[:, 7])))
    print("block us: mean %.1f median %.1f p90 %.1f p99 %.1f max %.1f" % (dur.mean(), np.median(dur), np.percentile(dur, 90), np.percentile(dur, 99), dur.max()))
    print("list A phase us: mean %.1f max %.1f" % (a[:, 2].mean() / 1e3, a[:, 2].max() / 1e3))
    print("warp imbalance: mean(max warp)/mean(mean warp) = %.2f" % (dur.mean() / warp.mean()))
    print("na: mean %.0f max %d   nb: mean %.0f max %d" % (a[:, 3].mean(), a[:, 3].max(), a[:, 4].mean(), a[:, 4].max()))
    for bit, name in ((1, "use_b"), (2, "no use_b"), (4, "use_a"), (8, "no use_a"), (0x700, "ball")):
        m = (a[:, 5] & bit) != 0
        print("  %-9s blocks %5d  mean us %.1f" % (name, m.sum(), dur[m].mean() if m.any() else 0.0))
    # when does each CTA finish
    ctas = {}
    for c, e in zip(a[:, 7], end):
        ctas[c] = max(ctas.get(c, 0.0), e)
    ends = np.array(sorted(ctas.values()))
    print("CTA finish us: min %.0f median %.0f max %.0f" % (ends.min(), np.median(ends), ends.max()))
    top = np.argsort(-dur)[:12]
    for i in top:
        print("  blk start %.0f dur %.0f na %d nb %d flags %x listA_us %.0f warps %s" % (start[i], dur[i], a[i, 3], a[i, 4], a[i, 5], a[i, 2] / 1e3, np.round(warp[i]).astype(int).tolist()))
    # correlation of duration with list sizes
    print("corr(dur, na) %.2f corr(dur, nb) %.2f" % (np.corrcoef(dur, a[:, 3])[0, 1], np.corrcoef(dur, a[:, 4])[0, 1]))


if __name__ == "__main__":
    main()
