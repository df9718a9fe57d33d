#!/bin/bash
# N-GPU visit of round 3: the sharded tests (incl. the skewed grow-and-repeat case), the weak-scaling
# bench at every N in "$2", and an N=1 bench line on HEAD (with the r03b capture matched).
# Usage: bash tools/gpu_r3_multi.sh <tag> "<Ns>"
TAG=${1:-r03c}; NS=${2:-"2"}; OUT=gpurun_out/$TAG; mkdir -p "$OUT"
nvidia-smi -L > "$OUT/gpus.txt" 2>&1
timeout 1500 python -m pytest tests/test_multi.py -m gpu -x -q > "$OUT/pytest_multi.log" 2>&1
echo "pytest multi exit $?" | tee -a "$OUT/pytest_multi.log"; tail -5 "$OUT/pytest_multi.log"
timeout 300 python bench.py --steps 20 --warmup 3 > "$OUT/bench_n1.json" 2> "$OUT/bench_n1.err"; echo "bench n1 exit $?"
for N in $NS; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus $N --steps 20 --warmup 3 --no-cpu > "$OUT/bench_n$N.json" 2> "$OUT/bench_n$N.err"
  echo "bench n$N exit $?"; python tools/benchsum.py < "$OUT/bench_n$N.json"
done
