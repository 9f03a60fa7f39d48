#!/bin/bash
# round 2, first GPU contact: GPU tests, bench line, reference arm
set -u
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/r2_pytest_gpu.log 2>&1; tail -15 $O/r2_pytest_gpu.log
python bench.py > $O/r2_bench.json 2> $O/r2_bench.err; tail -3 $O/r2_bench.err; cut -c1-1500 $O/r2_bench.json
