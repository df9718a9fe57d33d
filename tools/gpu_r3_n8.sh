#!/bin/bash
# The 8-GPU visit of round 3, must-have evidence first (VERDICT r2 "next" item 1): full-size configs 5 and 3
# on HEAD at N = 8, the weak-scaling bench at N = 8 and 4, the sharded tests at world 8, config 3 at N = 4.
TAG=${1:-r03e}; OUT=gpurun_out/$TAG; mkdir -p "$OUT"
nvidia-smi -L > "$OUT/gpus.txt" 2>&1
full() {  # cfg N timeout
  timeout $3 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port 2954$1 \
    tests/full_config.py --cfg $1 --window-bases 100e6 --out "$OUT/cfg$1_n$2.json" > "$OUT/cfg$1_n$2.log" 2>&1
  echo "== cfg $1 N=$2 exit $?"; grep "^{" "$OUT/cfg$1_n$2.log" | cut -c1-700
}
bench() {  # N timeout
  timeout $2 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 2955$1 \
    bench.py --gpus $1 --steps 20 --warmup 3 --no-cpu > "$OUT/bench_n$1.json" 2> "$OUT/bench_n$1.err"
  echo "== bench N=$1 exit $?"; python tools/benchsum.py < "$OUT/bench_n$1.json"
}
full 5 8 170
full 3 8 110
bench 8 110
bench 4 100
timeout 330 python -m pytest tests/test_multi.py -m gpu -x -q > "$OUT/pytest_multi.log" 2>&1
echo "pytest multi exit $?" | tee -a "$OUT/pytest_multi.log"; tail -3 "$OUT/pytest_multi.log"
full 3 4 100
