#!/bin/bash
# One 1-GPU visit of round 3: the allocation-free filter test alone in a fresh process, then
# tools/gpu_r2.sh (all -m gpu tests, bench with parity + cpu baseline, ncu launch list, ncu --set full
# on HEAD's default options).  Usage: bash tools/gpu_r3.sh <tag>
TAG=${1:-r03b}; OUT=gpurun_out/$TAG; mkdir -p "$OUT"
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu > "$OUT/zz_alone.log" 2>&1
echo "zz alone exit $?" | tee -a "$OUT/zz_alone.log"
bash tools/gpu_r2.sh "$TAG"
