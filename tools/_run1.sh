python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python tools/block_log.py 65536 2>&1 | head -16
python tools/block_log.py 16384 2>&1 | head -10
for wl in scaled_1k_1080p_d0 scaled_16k_1080p_d0 scaled_64k_1080p_d0; do python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['config']['workload'], d['ms_per_step'], d['e2e']['checksum'])"; done
