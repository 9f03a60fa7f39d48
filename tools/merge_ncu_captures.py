"""Folds gpurun_out/*.capture.json (tools/ncu_capture_json.py, written on the GPU box) into
profiles/ncu_captures.json and copies the matching summaries gpurun_out/<tag>.txt into profiles/."""
import glob, json, os, shutil
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
path = os.path.join(ROOT, "profiles", "ncu_captures.json")
cap = {"note": "figures bench.py may quote; each entry names the hash of the kernel sources it was captured on", "kernels": {}}
if os.path.exists(path):
    cap = json.load(open(path))
for f in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", "*.capture.json"))):
    frag = json.load(open(f))
    cap["kernels"][frag["workload"]] = frag
    txt = f.replace(".capture.json", ".txt")
    if os.path.exists(txt):
        shutil.copy(txt, os.path.join(ROOT, frag["file"]))
    print("merged", frag["workload"], frag["csrc_sha"], frag["file"])
json.dump(cap, open(path, "w"), indent=1)
