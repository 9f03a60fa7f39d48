# A/B of the bounce-ray grouping thresholds (environment knobs, same image for any value)
for k in 0.003 0.01; do for c in 0.94 0.9 0.85; do
echo "K=$k COS=$c"; RT_GROUP_K=$k RT_GROUP_COS=$c python tools/ab_modes.py 16384 1920 1080 2 0 | grep -o '"kernel_ms": [0-9.]*'; RT_GROUP_K=$k RT_GROUP_COS=$c python tools/ab_modes.py 1024 1920 1080 2 0 | grep -o '"kernel_ms": [0-9.]*'
done; done
RT_GROUP_COS=0.95 python tools/prof_scaled.py 16384 1920 1080 2 1 | tail -1
RT_GROUP_COS=0.95 python tools/prof_scaled.py 1024 1920 1080 2 1 | tail -1
