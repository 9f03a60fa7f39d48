"""Executed SASS footprint of the profiled kernel from an .ncu-rep (--import-source on): how many distinct
instructions ran, and how many bytes of code cover 50 / 90 / 99 % of the executed instructions.
python tools/ncu_footprint.py rep"""
import csv, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = None
ins = []
for r in rows:
    if r and "Instructions Executed" in r and hdr is None:
        hdr = r
        continue
    if hdr and len(r) == len(hdr):
        try:
            ins.append(int(r[hdr.index("Instructions Executed")] or 0))
        except ValueError:
            pass
tot = sum(ins)
ran = [x for x in ins if x > 0]
print("SASS instructions: %d (%.1f KB), executed at least once: %d (%.1f KB)" % (len(ins), len(ins) * 16 / 1024, len(ran), len(ran) * 16 / 1024))
ran.sort(reverse=True)
acc, k = 0, 0
marks = [0.5, 0.9, 0.99]
for i, x in enumerate(ran):
    acc += x
    while k < len(marks) and acc >= marks[k] * tot:
        print("  %2.0f %% of the executed instructions come from %5d instructions = %.1f KB" % (100 * marks[k], i + 1, (i + 1) * 16 / 1024))
        k += 1
