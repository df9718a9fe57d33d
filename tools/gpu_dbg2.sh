#!/bin/bash
# Time line of device-resident passes on N GPUs (OVL_DEBUG=1), optionally with bench options.
# Usage: bash tools/gpu_dbg2.sh <tag> "<list of N>" [--opt name=value ...]
TAG=${1:-r02d}; NS=${2:-2}; shift; shift
OUT=gpurun_out/$TAG; mkdir -p "$OUT"
for N in $NS; do
  OVL_DEBUG=1 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$N \
    bench.py --gpus $N --steps 6 --warmup 3 "$@" > "$OUT/dbg_n$N.log" 2> "$OUT/dbg_n$N.err"
  echo "== N=$N exit $?"; grep "^{" "$OUT/dbg_n$N.log" > "$OUT/bench_n$N.json"; python tools/benchsum.py < "$OUT/bench_n$N.json"
  # device-resident passes only (pack_end well under 1 ms), last ones
  grep -a "ovl rank" "$OUT/dbg_n$N.err" | grep -a "pack_end 0\." | tail -$((2*N)) | cut -c1-400
done
