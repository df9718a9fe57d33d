#!/bin/bash
# Option sweep on the bench workload.  Usage: bash tools/gpu_sweep.sh <tag> "<opt=value list, one bench run each; use , to join several>"
TAG=${1:-sw}; shift
OUT=gpurun_out/$TAG; mkdir -p "$OUT"
for o in $1; do
  args=""; for kv in ${o//,/ }; do args="$args --opt $kv"; done
  echo "== $o"
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu $args 2>&1 | tee -a "$OUT/sweep.jsonl" | python tools/benchsum.py
done
