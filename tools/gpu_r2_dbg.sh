#!/bin/bash
# Debug visit on N GPUs: all -m gpu tests, bench at 1 and N GPUs, and a short N-GPU bench with the
# OVL_DEBUG=1 stream time line.  Usage: bash tools/gpu_r2_dbg.sh <tag> <N>
TAG=${1:-r02d}; N=${2:-2}
OUT=gpurun_out/$TAG; mkdir -p "$OUT"
timeout 1500 python -m pytest tests -m gpu -q -x > "$OUT/pytest_gpu.log" 2>&1; echo "pytest exit $?" >> "$OUT/pytest_gpu.log"; tail -12 "$OUT/pytest_gpu.log"
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 3 --no-cpu > "$OUT/bench_n1.json" 2> "$OUT/bench_n1.err"
echo "== N=1 exit $?"; python tools/benchsum.py < "$OUT/bench_n1.json"; tail -3 "$OUT/bench_n1.err"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
  bench.py --gpus $N --steps 20 --warmup 3 > "$OUT/bench_n$N.log" 2> "$OUT/bench_n$N.err"
grep "^{" "$OUT/bench_n$N.log" > "$OUT/bench_n$N.json"
echo "== N=$N exit $?"; python tools/benchsum.py < "$OUT/bench_n$N.json"; grep -v "OMP_NUM\|^\*\*\*\|^$" "$OUT/bench_n$N.err" | tail -6
OVL_DEBUG=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 \
  bench.py --gpus $N --steps 3 --warmup 3 --no-parity > "$OUT/dbg_n$N.log" 2> "$OUT/dbg_n$N.err"
grep "ovl rank" "$OUT/dbg_n$N.err" | tail -$((4*N))
OVL_DEBUG=1 timeout 300 python bench.py --gpus 1 --steps 3 --warmup 3 --no-parity --no-cpu > "$OUT/dbg_n1.log" 2> "$OUT/dbg_n1.err"
grep "ovl rank" "$OUT/dbg_n1.err" | tail -3
