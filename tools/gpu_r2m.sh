#!/bin/bash
# N-GPU visit: all -m gpu tests (1-GPU and sharded), then bench with parity + device time lines.
# Usage: bash tools/gpu_r2m.sh <tag> "<list of N>"
TAG=${1:-r02m}; NS=${2:-"1 2"}
OUT=gpurun_out/$TAG; mkdir -p "$OUT"
timeout 1500 python -m pytest tests -m gpu -q -x > "$OUT/pytest_gpu.log" 2>&1; echo "pytest exit $?" >> "$OUT/pytest_gpu.log"; tail -6 "$OUT/pytest_gpu.log"
bash tools/gpu_dbg2.sh $TAG "$NS"
