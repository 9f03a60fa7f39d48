#!/bin/bash
# Round-2 8-GPU set: parity of both gathers, then the bench line at N = 2, 4, 8 (default workload: carries
# bands_gather, bands_gather_4k_spp16, batch256), once more at N = 8 without the NUMA memory policy (A/B).
set -u
O=gpurun_out; mkdir -p $O
run() { n=$1; shift; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) "$@" 2>>$O/r2_mgpu.err | grep -E '^\{'; }
nvidia-smi topo -m > $O/r2_topo.txt 2>&1; numactl -H >> $O/r2_topo.txt 2>&1; ls /sys/devices/system/node >> $O/r2_topo.txt 2>&1; nproc >> $O/r2_topo.txt
run 8 tools/mgpu_check.py > $O/r2_mgpu8_check.json
for n in 2 4 8; do run $n bench.py --gpus $n --steps 50 --warmup 3 >> $O/r2_scale_default.jsonl; done
RT_NO_NUMA_MEMORY=1 run 8 bench.py --gpus 8 --steps 50 --warmup 3 --no-legs > $O/r2_scale8_no_numa_policy.json
run 8 bench.py --gpus 8 --impl reference --steps 2 --warmup 0 > $O/r2_reference_arm_under_torchrun8.json
python - <<PY
import json
for l in open("$O/r2_scale_default.jsonl"):
    d=json.loads(l)
    print("N", d["n_gpus"], "ms", round(d["ms_per_step"],4), "e2e ms", round(d["e2e"]["ms_per_step"],4), d["host_binding"][:150])
    for k in ("bands_gather","bands_gather_4k_spp16"):
        b=d.get(k,{})
        for m in ("nccl_gather","peer_store","peer_store_feedback_bands"):
            if m in b: print("  ",k,m, {x:(round(b[m][x],4) if isinstance(b[m].get(x),float) else b[m].get(x)) for x in ("ms_per_frame","speedup_vs_single_gpu","byte_identical_to_single_gpu","imbalance_max_over_mean","unavailable")})
    print("   batch256", {k:v for k,v in d.get("batch256",{}).items() if k in ("ms_per_batch","value","first_frame_equals_single_call")}, d.get("batch256",{}).get("e2e"))
d=json.load(open("$O/r2_scale8_no_numa_policy.json")); print("N 8 no policy: e2e ms", d["e2e"]["ms_per_step"], d["host_binding"][-80:])
d=json.load(open("$O/r2_reference_arm_under_torchrun8.json")); print("reference arm under torchrun 8:", d["value"], d["cpu_baseline"]["cores"])
PY
