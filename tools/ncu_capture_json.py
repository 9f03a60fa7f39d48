"""Records what bench.py reports from an `ncu --set full` capture (DRAM traffic per launch, issue-slot use) together
with the hash of the kernel sources the capture was taken on:  python tools/ncu_capture_json.py rep workload tag
-> gpurun_out/<tag>.capture.json; tools/merge_ncu_captures.py folds the fragments into profiles/ncu_captures.json.
bench.py prints the figures only while the hash still matches the sources it runs."""
import csv, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

rep, workload, tag = sys.argv[1], sys.argv[2], sys.argv[3]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2]


def val(name, scale_units=True):
    i = hdr.index(name)
    v = float(data[i].replace(",", ""))
    u = units[i].lower()
    if scale_units:
        v *= {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "us": 1e-3, "ns": 1e-6, "ms": 1.0, "msecond": 1.0,
              "usecond": 1e-3, "nsecond": 1e-6, "second": 1e3}.get(u, 1)
    return v


frag = {"workload": workload, "csrc_sha": bench.csrc_sha(), "file": "profiles/%s.txt" % tag,
        "kernel": data[hdr.index("Kernel Name")],
        "dram_bytes_read": int(val("dram__bytes_read.sum")), "dram_bytes_write": int(val("dram__bytes_write.sum")),
        "issue_active": val("smsp__issue_active.avg.pct_of_peak_sustained_active") / 100.0,
        "duration_ms_under_ncu": val("gpu__time_duration.sum"),
        "registers": int(val("launch__registers_per_thread", False)),
        "inst_executed": int(val("smsp__inst_executed.sum", False))}
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
with open(os.path.join(ROOT, "gpurun_out", tag + ".capture.json"), "w") as f:
    json.dump(frag, f, indent=1)
print(json.dumps(frag))
