#!/bin/bash
# Scaling bench only (no tests).  Usage: bash tools/gpu_scale_only.sh <tag> "<list of N>" [bench args]
TAG=${1:-s}; NS=${2:-"2"}; shift; shift
OUT=gpurun_out/$TAG; mkdir -p "$OUT"
for n in $NS; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n \
    bench.py --gpus $n --steps 20 --warmup 3 "$@" > "$OUT/bench_n$n.log" 2> "$OUT/bench_n$n.err"
  grep "^{" "$OUT/bench_n$n.log" > "$OUT/bench_n$n.json"
  echo "== N=$n exit $?"; python tools/benchsum.py < "$OUT/bench_n$n.json"; grep -v "OMP_NUM\|^\*\*\*" "$OUT/bench_n$n.err" | tail -3
done
