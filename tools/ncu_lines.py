"""Per-source-line executed instructions and stall samples from an .ncu-rep
(needs -lineinfo and --import-source on).  python tools/ncu_lines.py rep [top]"""
import csv, subprocess, sys
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = None
lines = []
for r in rows:
    if r and r[0] == "Line No":
        hdr = r
        continue
    if hdr and len(r) == len(hdr) and r[2] == "-":      # source-line summary rows
        lines.append(r)
ix = {h: i for i, h in enumerate(hdr)}
iex, ismp = hdr.index("Instructions Executed"), hdr.index("# Samples")
tot_i = sum(int(r[iex] or 0) for r in lines)
tot_s = sum(int(r[ismp] or 0) for r in lines)
print("total inst %.4g samples %d" % (tot_i, tot_s))
lines.sort(key=lambda r: -int(r[ismp] or 0))
for r in lines[:top]:
    print("%5.1f%% smp %5.1f%% inst  L%-4s %s" % (100.0 * int(r[ismp] or 0) / tot_s, 100.0 * int(r[iex] or 0) / tot_i,
                                                 r[0], r[1].strip()[:110]))
