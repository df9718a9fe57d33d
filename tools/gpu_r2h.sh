#!/bin/bash
# 1-GPU visit: all GPU tests (default kernels), the same parity tests with the bulk-async filter,
# bench A/B of the two filter kernels, then compute-sanitizer.  Usage: bash tools/gpu_r2h.sh <tag>
TAG=${1:-r02h}; OUT=gpurun_out/$TAG; mkdir -p "$OUT"
timeout 900 python -m pytest tests -m gpu -x -q > "$OUT/pytest_gpu.log" 2>&1; echo "pytest exit $?" >> "$OUT/pytest_gpu.log"; tail -6 "$OUT/pytest_gpu.log"
OVL_SCAN_BULK=1 timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > "$OUT/pytest_bulk.log" 2>&1; echo "pytest bulk exit $?" >> "$OUT/pytest_bulk.log"; tail -6 "$OUT/pytest_bulk.log"
for o in scan_bulk=0 scan_bulk=1; do
  echo "== $o"
  timeout 200 python bench.py --steps 20 --warmup 3 --no-cpu --opt $o > "$OUT/bench_$o.json" 2> "$OUT/bench_$o.err"
  echo "exit $?"; python tools/benchsum.py < "$OUT/bench_$o.json"; tail -3 "$OUT/bench_$o.err"
done
bash tools/gpu_sanitize.sh $TAG
