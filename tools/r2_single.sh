#!/bin/bash
# Round-2 single-GPU measurement set (run under gpurun): GPU tests, bench line (with its scaled legs) and reference arm,
# depth-2 scaled workloads, the ncu launch list of the default bench command (after it exited 0 without ncu), and the
# `ncu --set full` captures bench.py quotes (profiles/ncu_captures.json).  Files land in gpurun_out/r2_*.
set -u
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -q > $O/r2_pytest_gpu.log 2>&1; tail -2 $O/r2_pytest_gpu.log
python bench.py > $O/r2_bench_builtin_1080p_d4.json 2> $O/r2_bench.err; tail -2 $O/r2_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/r2_bench_reference_arm.json 2>> $O/r2_bench.err
rm -f $O/r2_bench_other_workloads.jsonl
for wl in scaled_1k_1080p_d2 scaled_16k_1080p_d2 scaled_4k_1080p_d0 scaled_16k_1080p_d0 scaled_64k_1080p_d0 builtin_720p_d0 builtin_720p_d4 builtin_1080p_d0 builtin_4k_spp16_d4 batch256_1024_d4; do
  python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu-baseline --no-legs >> $O/r2_bench_other_workloads.jsonl 2>> $O/r2_bench.err
done
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-legs > $O/r2_plain.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/r2_launches_builtin_1080p_d4.csv \
      python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-legs > $O/r2_ncu_launches.log 2>&1
bash tools/run_ncu.sh r2_ncu_full_builtin_1080p_d4 5 1920 1080 4 default builtin_1080p_d4 > /dev/null
bash tools/run_ncu.sh r2_ncu_full_scaled1k_d0 1024 1920 1080 0 default scaled_1k_1080p_d0 > /dev/null
bash tools/run_ncu.sh r2_ncu_full_scaled64k_d0 65536 1920 1080 0 default scaled_64k_1080p_d0 > /dev/null
bash tools/run_ncu.sh r2_ncu_full_scaled1k_d2 1024 1920 1080 2 > /dev/null
bash tools/run_ncu.sh r2_ncu_full_scaled16k_d2_960x540 16384 960 540 2 > /dev/null
rm -f $O/*_sass.csv $O/*_all_lines.txt
ls $O | head -40
