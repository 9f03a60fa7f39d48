#!/bin/bash
# A/B of the object groups (no_cull bit 8) and of the block / per-warp kernels with them: median kernel ms, frames compared
set -u
O=gpurun_out; mkdir -p $O
{
for n in 2048 4096 16384; do echo "== d2 n=$n modes 0 256 8 16 (default, no groups, per-warp, block)"; python tools/ab_modes.py $n 1920 1080 2 0 256 8 16; done
for n in 4096 16384 65536; do echo "== d0 n=$n modes 0 256 8"; python tools/ab_modes.py $n 1920 1080 0 0 256 8; done
echo "== RT_GROUPS_MIN=256: d0 n=1024 modes 0 256; d2 n=1024 modes 0 256 8 16"
RT_GROUPS_MIN=256 python tools/ab_modes.py 1024 1920 1080 0 0 256
RT_GROUPS_MIN=256 python tools/ab_modes.py 1024 1920 1080 2 0 256 8 16
} 2>&1 | tee $O/ab_groups.txt
