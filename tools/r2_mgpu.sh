#!/bin/bash
# multi-GPU check + bench line at N ranks: tools/r2_mgpu.sh N [tag]
set -u
N=$1; T=${2:-mg$1}; O=gpurun_out; mkdir -p $O
run() { timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) "$@"; }
run tools/mgpu_check.py > $O/${T}_check.json 2> $O/${T}_check.err; tail -c 1500 $O/${T}_check.json; tail -3 $O/${T}_check.err
run bench.py --gpus $N --steps 50 --warmup 3 > $O/${T}_bench.json 2> $O/${T}_bench.err; tail -3 $O/${T}_bench.err
python - <<PY
import json
d=[json.loads(l) for l in open("$O/${T}_bench.json") if l.startswith("{")][0]
print("ms_per_step", d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"])
for k in ("bands_gather","bands_gather_4k_spp16"):
    b=d.get(k,{})
    print(k, "single", b.get("single_gpu_ms"))
    for m in ("nccl_gather","peer_store","peer_store_feedback_bands"):
        if m in b: print("  ",m, {x:b[m].get(x) for x in ("ms_per_frame","speedup_vs_single_gpu","byte_identical_to_single_gpu","imbalance_max_over_mean","completion_and_launch_ms","wait_timed_out","unavailable")})
print("batch256", {k:v for k,v in d.get("batch256",{}).items() if k in ("ms_per_batch","value","e2e","first_frame_equals_single_call")})
PY
