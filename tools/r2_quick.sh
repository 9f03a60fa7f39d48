#!/bin/bash
# quick GPU check of a kernel change: smoke + the fast subset of the parity tests + the bench line without legs
set -u
O=gpurun_out; mkdir -p $O
T=${1:-quick}
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "builtin_scene_vs_oracle or golden or scaled_scene_small or bounding_sphere_test_changes or edge_case or tie or ragged or inside_box" > $O/${T}_pytest.log 2>&1; tail -3 $O/${T}_pytest.log
python bench.py --no-legs --no-cpu-baseline --steps 50 > $O/${T}_bench.json 2> $O/${T}_bench.err; tail -3 $O/${T}_bench.err
python - <<PY
import json
d=json.load(open("$O/${T}_bench.json"))
print("ms_per_step", d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], "frac", d["roofline"]["frac"], "checksum", d["e2e"]["checksum"])
PY
