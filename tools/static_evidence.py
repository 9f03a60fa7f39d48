"""Static evidence of the built kernels (no GPU needed): copies the ptxas -v logs of the last `make` into profiles/,
writes a one-line-per-kernel summary of them, and counts in `cuobjdump -sass` of rt_kernel_fast.o the opcodes that
prove the data-movement path.  python tools/static_evidence.py [prefix, default r2]"""
import os
import re
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "openglraytracer_b200", "csrc", "build")
PROF = os.path.join(ROOT, "profiles")
pre = sys.argv[1] if len(sys.argv) > 1 else "r2"


def demangle(names):
    out = subprocess.run(["c++filt"] + names, capture_output=True, text=True).stdout.split("\n")
    return dict(zip(names, out))


def ptxas_summary(log):
    rows, cur = [], None
    for line in open(log):
        m = re.search(r"Compiling entry function '(\S+)'", line)
        if m:
            cur = {"name": m.group(1)}
            rows.append(cur)
            continue
        if cur is None:
            continue
        m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", line)
        if m and "stack" not in cur:
            cur["stack"], cur["st"], cur["ld"] = (int(x) for x in m.groups())
        m = re.search(r"Used (\d+) registers", line)
        if m and "regs" not in cur:
            cur["regs"] = int(m.group(1))
    names = demangle([r["name"] for r in rows])
    return ["%-75s regs %3d  stack %5d B  spill stores %5d B  loads %5d B" %
            (names[r["name"]], r.get("regs", -1), r.get("stack", 0), r.get("st", 0), r.get("ld", 0)) for r in rows]


def sass_counts(obj):
    sass = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
    ops = ["UBLKCP", "SYNCS", "LDGSTS", "LDS.128", "STG.E.128", "FFMA", "MUFU", "LDL", "STL"]
    rows, cur, body = [], None, {}
    for line in sass.split("\n"):
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            rows.append(cur)
            body[cur] = []
            continue
        if cur and re.search(r"/\*[0-9a-f]{4,}\*/\s+\S", line):
            body[cur].append(line)
    names = demangle(rows)
    out = []
    for r in rows:
        if "Kernel" not in names[r] and "kernel" not in names[r]:
            continue
        text = body[r]
        cnt = {o: sum(1 for l in text if re.search(r"\s" + re.escape(o) + r"[\s.]", l)) for o in ops}
        out.append("%-75s %6d instr  " % (names[r], len(text)) + "  ".join("%s %d" % (o, cnt[o]) for o in ops))
    # excerpt: the streaming tile issue / wait of the first kernel that has a UBLKCP
    ex = []
    for r in rows:
        idx = [i for i, l in enumerate(body[r]) if "UBLKCP" in l]
        if idx:
            i0 = max(0, idx[0] - 8)
            ex = ["# excerpt: streaming tile issue + wait in %s (mbarrier.arrive.expect_tx -> SYNCS.ARRIVE.TRANS64, "
                  "cp.async.bulk -> UBLKCP, try_wait -> SYNCS.PHASECHK)" % names[r]] + body[r][i0:idx[0] + 6]
            ph = [i for i, l in enumerate(body[r]) if "SYNCS.PHASECHK" in l]
            if ph:
                ex += ["        ..."] + body[r][max(0, ph[0] - 2):ph[0] + 3]
            break
    return out, ex


for kind in ("fast", "strict"):
    shutil.copy(os.path.join(BUILD, "ptxas_%s.log" % kind), os.path.join(PROF, "%s_ptxas_%s.log" % (pre, kind)))
with open(os.path.join(PROF, "%s_ptxas_fast_summary.txt" % pre), "w") as f:
    f.write("# ptxas -v summary of rt_kernel_fast.cu (sm_100a), one line per kernel entry; full log: %s_ptxas_fast.log\n" % pre)
    f.write("\n".join(ptxas_summary(os.path.join(BUILD, "ptxas_fast.log"))) + "\n")
rows, ex = sass_counts(os.path.join(BUILD, "rt_kernel_fast.o"))
with open(os.path.join(PROF, "%s_sass_evidence.txt" % pre), "w") as f:
    f.write("# cuobjdump -sass of rt_kernel_fast.o (sm_100a): per kernel, instruction counts of the opcodes that prove the "
            "data-movement path\n# UBLKCP = cp.async.bulk (TMA bulk copy), SYNCS = mbarrier ops, LDGSTS = cp.async, LDS.128 = "
            "16-byte shared loads, STG.E.128 = 16-byte framebuffer stores\n\n")
    f.write("\n".join(rows) + "\n\n" + "\n".join(ex) + "\n")
print("\n".join(rows[:6]))
