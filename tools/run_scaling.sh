#!/bin/bash
# Multi-GPU measurement set (run under gpurun --gpus 8): band/peer-store parity, then the 4K spp16
# frame (config 3), the 256-pose batch (config 5) and the default workload at N = 2, 4, 8.
# Usage: tools/run_scaling.sh [prefix]   (files land in gpurun_out/<prefix>_*.jsonl, default r1)
set -u
P=${1:-r1}
mkdir -p gpurun_out
run() { n=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) "$@" 2>gpurun_out/${P}_last.err | grep -E '^\{' ; }
run 8 tools/mgpu_check.py > gpurun_out/${P}_mgpu8_check.json
for n in 2 4 8; do
  run $n bench.py --gpus $n --workload builtin_4k_spp16_d4 --steps 10 --warmup 3 --no-cpu-baseline >> gpurun_out/${P}_scale_4k_spp16.jsonl
  run $n bench.py --gpus $n --workload batch256_1024_d4 --steps 5 --warmup 3 --no-cpu-baseline >> gpurun_out/${P}_scale_batch256.jsonl
  run $n bench.py --gpus $n --steps 30 --warmup 3 --no-cpu-baseline >> gpurun_out/${P}_scale_default.jsonl
done
run 8 bench.py --gpus 8 --workload scaled_64k_1080p_d0 --steps 5 --warmup 3 --no-cpu-baseline >> gpurun_out/${P}_scale_scaled64k.jsonl
