#!/bin/bash
# scaled-scene check: device-prologue parity tests + the bench line with its scaled legs
set -u
O=gpurun_out; mkdir -p $O; T=${1:-sc}
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "device_prologue or scaled or bounding or large_scene or caller_stream" > $O/${T}_pytest.log 2>&1; tail -3 $O/${T}_pytest.log
python bench.py --no-cpu-baseline --steps 20 > $O/${T}_bench.json 2> $O/${T}_bench.err; tail -3 $O/${T}_bench.err
python - <<PY
import json
d=[json.loads(l) for l in open("$O/${T}_bench.json") if l.startswith("{")][0]
print("default ms", d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"])
for k,v in d["scaled"].items():
    if k=="note": continue
    nc = v.get("every_slab_test_no_cull_1")
    print(k, "shipped %.4f ms frac %.3f" % (v["shipped"]["kernel_ms"], v["shipped"]["frac_of_fp32_peak"]),
          ("nocull %.3f ms frac %.3f" % (nc["kernel_ms"], nc["frac_of_fp32_peak"])) if nc else "", "e2e %.4f ms (pageable %.4f)" % (v["e2e"]["ms_per_step"], v["e2e"]["ms_per_step_pageable_objects"]), v["e2e"]["frame_equals_device_frame"])
PY
