#!/bin/bash
# N-GPU visit: the sharded tests (world = all GPUs of the box), then bench with parity + time lines.
# Usage: bash tools/gpu_r2p.sh <tag> "<list of N>"
TAG=${1:-r02p}; NS=${2:-"2 4"}
OUT=gpurun_out/$TAG; mkdir -p "$OUT"
timeout 1200 python -m pytest tests/test_multi.py -m gpu -q -x > "$OUT/pytest_multi.log" 2>&1; echo "pytest exit $?" >> "$OUT/pytest_multi.log"; tail -5 "$OUT/pytest_multi.log"
bash tools/gpu_dbg2.sh $TAG "$NS"
