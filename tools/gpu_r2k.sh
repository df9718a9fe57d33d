#!/bin/bash
# 1-GPU visit: option sweep, then the full-size configs 1, 2, 4 (and 5 when the host has the memory)
# through tests/full_config.py.  Usage: bash tools/gpu_r2k.sh <tag>
TAG=${1:-r02k}; OUT=gpurun_out/$TAG; mkdir -p "$OUT"
bash tools/gpu_r2i.sh $TAG "k2_blocks_per_sm=4,pack_blocks_per_sm=8 k2_blocks_per_sm=3,pack_blocks_per_sm=8 k2_blocks_per_sm=6,pack_blocks_per_sm=8 k2_blocks_per_sm=4,pack_blocks_per_sm=6 k2_blocks_per_sm=4,pack_blocks_per_sm=10 anchor_stride=64,k2_blocks_per_sm=8,pack_blocks_per_sm=8" sweep-only
bash tools/gpu_full.sh $TAG 1 "1 2 4" 40e6
AVAIL=$(free -g | awk '/^Mem:/{print $7}')
echo "host MemAvailable ${AVAIL} GB"
if [ "${AVAIL:-0}" -ge 120 ]; then bash tools/gpu_full.sh $TAG 1 "5" 100e6; else echo "config 5 on one GPU skipped: host memory"; fi
