#!/bin/bash
# Multi-GPU visit (gpurun --gpus N): sharded parity tests, then the weak-scaling bench at 1/2/4/8.
# Usage: bash tools/gpu_scale.sh <tag> "<list of N>"
TAG=${1:-s}; NS=${2:-"1 2 4 8"}
OUT=gpurun_out/$TAG; mkdir -p "$OUT"
nvidia-smi -L > "$OUT/gpus.txt"
timeout 1200 python -m pytest tests/test_multi.py -m gpu -x -q > "$OUT/pytest_multi.log" 2>&1; echo "pytest exit $?" >> "$OUT/pytest_multi.log"; tail -4 "$OUT/pytest_multi.log"
for n in $NS; do
  if [ "$n" = 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 20 --warmup 3 --no-cpu > "$OUT/bench_n$n.json" 2> "$OUT/bench_n$n.err"
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n \
      bench.py --gpus $n --steps 20 --warmup 3 > "$OUT/bench_n$n.log" 2> "$OUT/bench_n$n.err"
    grep "^{" "$OUT/bench_n$n.log" > "$OUT/bench_n$n.json"
  fi
  echo "== N=$n exit $?"; python tools/benchsum.py < "$OUT/bench_n$n.json"; tail -3 "$OUT/bench_n$n.err"
done
