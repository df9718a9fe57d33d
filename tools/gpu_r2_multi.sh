#!/bin/bash
# Multi-GPU visit of round 2 (gpurun --gpus N): sharded parity tests, CLI --gpus test, weak-scaling bench
# with the parity field.  Usage: bash tools/gpu_r2_multi.sh <tag> "<list of N>" [skip-tests] [bench args]
TAG=${1:-r02s}; NS=${2:-"2"}; SKIP=${3:-}; shift; shift; shift
OUT=gpurun_out/$TAG; mkdir -p "$OUT"
nvidia-smi -L > "$OUT/gpus.txt"; nvidia-smi topo -m >> "$OUT/gpus.txt" 2>&1
if [ "$SKIP" != "skip-tests" ]; then
  timeout 1500 python -m pytest tests/test_multi.py -m gpu -q > "$OUT/pytest_multi.log" 2>&1; echo "pytest exit $?" >> "$OUT/pytest_multi.log"; tail -25 "$OUT/pytest_multi.log"
fi
for n in $NS; do
  if [ "$n" = 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 20 --warmup 3 --no-cpu "$@" > "$OUT/bench_n$n.json" 2> "$OUT/bench_n$n.err"
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n \
      bench.py --gpus $n --steps 20 --warmup 3 "$@" > "$OUT/bench_n$n.log" 2> "$OUT/bench_n$n.err"
    grep "^{" "$OUT/bench_n$n.log" > "$OUT/bench_n$n.json"
  fi
  echo "== N=$n exit $?"; python tools/benchsum.py < "$OUT/bench_n$n.json"; grep -v "OMP_NUM\|^\*\*\*\|^$" "$OUT/bench_n$n.err" | tail -6
done
