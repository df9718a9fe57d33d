#!/bin/bash
# One 1-GPU visit of round 2: parity tests, bench (with parity + cpu baseline), and -- only when both
# exited 0 -- the ncu launch list and the ncu --set full capture of the hot kernels, each behind a
# plain run of the very same command.  Usage: bash tools/gpu_r2.sh <tag> [sweep]
set -u
TAG=${1:-r02a}
OUT=gpurun_out/$TAG
mkdir -p "$OUT"
nvidia-smi > "$OUT/nvidia_smi.txt" 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q > "$OUT/pytest_gpu.log" 2>&1; T=$?
echo "pytest exit $T" >> "$OUT/pytest_gpu.log"; tail -15 "$OUT/pytest_gpu.log"
timeout 600 python bench.py --steps 20 --warmup 3 > "$OUT/bench.json" 2> "$OUT/bench.err"; B=$?
echo "bench exit $B"; python tools/benchsum.py < "$OUT/bench.json"; tail -5 "$OUT/bench.err"
if [ "${2:-}" = "sweep" ]; then
  for o in anchor_stride=64 anchor_stride=16 filter_kb=4096 filter_kb=32768; do
    echo "== $o"
    timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-parity --opt $o 2>&1 | tee -a "$OUT/sweep.jsonl" | python tools/benchsum.py
  done
fi
if [ $T -eq 0 ] && [ $B -eq 0 ]; then
  CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-parity"
  timeout 300 $CMD > "$OUT/plain.log" 2>&1 &&
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
    --log-file "$OUT/launches.csv" $CMD > "$OUT/ncu_list.log" 2>&1
  echo "ncu list exit $?"
  timeout 300 $CMD > "$OUT/plain2.log" 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on \
    --kernel-name regex:'k1_pack|k3_filter|k3_resolve|k4a_head|k4b_tail|k2_insert|k7_' \
    --launch-skip 55 --launch-count 11 -o "$OUT/full" -f $CMD > "$OUT/ncu_full.log" 2>&1
  echo "ncu full exit $?"
fi
ls -la "$OUT"
