#!/bin/bash
# 1-GPU visit: option sweep (one bench line per variant), then the full tools/gpu_r2.sh capture
# (tests, bench with parity + cpu baseline, ncu launch list, ncu --set full).  Usage: bash tools/gpu_r2i.sh <tag> "<opts>"
TAG=${1:-r02i}; OUT=gpurun_out/$TAG; mkdir -p "$OUT"
OPTS=${2:-"k2_blocks_per_sm=1 k2_blocks_per_sm=2 k2_blocks_per_sm=4 k2_blocks_per_sm=8 anchor_stride=64 scan_bulk=1 pack_blocks_per_sm=5 pack_blocks_per_sm=8"}
echo "== default"
timeout 200 python bench.py --steps 20 --warmup 3 --no-cpu --no-parity 2>&1 | tee -a "$OUT/sweep.jsonl" | python tools/benchsum.py
for o in $OPTS; do
  echo "== $o"
  timeout 200 python bench.py --steps 20 --warmup 3 --no-cpu --no-parity $(for x in ${o//,/ }; do echo --opt $x; done) 2>&1 | tee -a "$OUT/sweep.jsonl" | python tools/benchsum.py
done
[ "${3:-}" = "sweep-only" ] || bash tools/gpu_r2.sh $TAG
