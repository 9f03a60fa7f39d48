# A/B of the bounce-ray grouping thresholds (environment knobs, same image for any value): 16k objects depth 2 (per-warp
# kernel, object groups on) and 1k objects depth 2 (block kernel)
for k in 0.003 0.01 0.03; do for c in 0.98 0.96 0.94 0.9; do
echo "K=$k COS=$c"; RT_GROUP_K=$k RT_GROUP_COS=$c python tools/ab_modes.py 16384 1920 1080 2 0 | grep -o '"kernel_ms": [0-9.]*'; RT_GROUP_K=$k RT_GROUP_COS=$c python tools/ab_modes.py 1024 1920 1080 2 0 | grep -o '"kernel_ms": [0-9.]*'
done; done
