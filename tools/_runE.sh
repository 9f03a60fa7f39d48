for n in 1024 4096 16384; do for m in default noblock; do echo -n "$n d2 1080p $m: "; python tools/prof_scaled.py $n 1920 1080 2 3 $m | tail -2 | python -c "import sys,json; print([round(json.loads(l)['kernel_ms'],2) for l in sys.stdin])"; done; done
for n in 1024 16384; do for m in default noblock; do echo -n "$n d1 1080p $m: "; python tools/prof_scaled.py $n 1920 1080 1 3 $m | tail -2 | python -c "import sys,json; print([round(json.loads(l)['kernel_ms'],2) for l in sys.stdin])"; done; done
python tools/check_modes.py 3 0 | cut -c1-120
python tools/stress_block.py 5 | grep -v "bad reps 0"
python tools/sanity_small.py 2>&1 | tail -3
