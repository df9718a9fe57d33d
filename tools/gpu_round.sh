#!/bin/bash
# One GPU-box visit: parity tests, bench, ncu launch list, ncu --set full of the hot kernels.
# Usage (from the repo root, under gpurun):  bash tools/gpu_round.sh <tag> [skip-tests]
set -u
TAG=${1:-rXX}
OUT=gpurun_out/$TAG
mkdir -p "$OUT"
nvidia-smi > "$OUT/nvidia_smi.txt" 2>&1
if [ "${2:-}" != "skip-tests" ]; then
  timeout 900 python -m pytest tests -m gpu -x -q > "$OUT/pytest_gpu.log" 2>&1
  echo "pytest exit $?" >> "$OUT/pytest_gpu.log"
  tail -5 "$OUT/pytest_gpu.log"
fi
timeout 600 python bench.py --steps 20 --warmup 3 > "$OUT/bench.json" 2> "$OUT/bench.err"
echo "bench exit $?"
tail -c 3000 "$OUT/bench.json"
tail -5 "$OUT/bench.err"
# launch list (cold-cache, serialised: compare shares only)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
  --log-file "$OUT/launches.csv" python bench.py --steps 2 --warmup 3 --no-cpu > "$OUT/ncu_list.log" 2>&1
echo "ncu list exit $?"
# full capture of our kernels, one launch each after warm-up (skip the first 3 of each)
timeout 900 ncu --set full --clock-control none --import-source on \
  --kernel-name regex:'k1_pack|k3_filter|k3_resolve|k4a_head|k4b_tail|k2_insert|k7_' \
  --launch-skip 55 --launch-count 11 -o "$OUT/full" -f \
  python bench.py --steps 2 --warmup 3 --no-cpu > "$OUT/ncu_full.log" 2>&1
echo "ncu full exit $?"
ls -la "$OUT"
