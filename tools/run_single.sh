#!/bin/bash
# Single-GPU measurement set (run under gpurun): GPU tests, the bench line and the reference arm, the
# scaled scenes with and without the bounding tests, configs 3 and 5, then the ncu launch list of the
# default bench command (only after the same command exited 0 without ncu).
# Usage: tools/run_single.sh [prefix]    (files land in gpurun_out/<prefix>_*)
set -u
P=${1:-r1}
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -q > $O/${P}_pytest_gpu.log 2>&1; tail -2 $O/${P}_pytest_gpu.log
python bench.py > $O/${P}_bench.json 2> $O/${P}_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/${P}_bench_ref.json 2>&1
for wl in scaled_1k_1080p_d0 scaled_4k_1080p_d0 scaled_16k_1080p_d0 scaled_64k_1080p_d0 scaled_1k_1080p_d2 scaled_16k_1080p_d2; do
  python bench.py --workload $wl --steps 10 --warmup 3 --cpu-seconds 6 >> $O/${P}_bench_scaled.jsonl 2>> $O/${P}_bench.err
done
for n in 1024 65536; do python tools/prof_scaled.py $n 1920 1080 0 2 nocull | tail -1 >> $O/${P}_scaled_nocull.jsonl; done
# A/B of the switches (same image): per-ray bounding tests only, per-warp packets, block kernel without fan planes, bounce rays ungrouped
for n in 1024 16384 65536; do for m in nopacket noblock nofan; do python tools/prof_scaled.py $n 1920 1080 0 2 $m | tail -1 >> $O/${P}_scaled_ab.jsonl; done; done
for n in 1024 16384; do for m in default nogroup; do python tools/prof_scaled.py $n 960 540 2 2 $m | tail -1 >> $O/${P}_scaled_ab.jsonl; done; done
python bench.py --workload batch256_1024_d4 --steps 5 --warmup 3 --no-cpu-baseline > $O/${P}_bench_batch256.json 2>> $O/${P}_bench.err
python bench.py --workload builtin_4k_spp16_d4 --steps 5 --warmup 3 --no-cpu-baseline > $O/${P}_bench_4k_spp16.json 2>> $O/${P}_bench.err
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/${P}_plain.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/${P}_launches.csv \
      python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/${P}_ncu_launches.log 2>&1
cat $O/${P}_bench.json
