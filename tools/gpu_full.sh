#!/bin/bash
# Full-size BASELINE configs through tests/full_config.py.  Usage: bash tools/gpu_full.sh <tag> <ngpus> "<cfg list>" [window-bases]
TAG=${1:-full}; N=${2:-1}; CFGS=${3:-"1 2 4"}; WIN=${4:-60e6}
OUT=gpurun_out/$TAG; mkdir -p "$OUT"
free -g | head -2 > "$OUT/host_mem.txt"; nproc >> "$OUT/host_mem.txt"
for c in $CFGS; do
  timeout ${FULL_TIMEOUT:-1500} python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$c \
    tests/full_config.py --cfg $c --window-bases $WIN --out "$OUT/cfg${c}_n$N.json" > "$OUT/cfg${c}_n$N.log" 2>&1
  echo "== cfg $c N=$N exit $?"; grep "^{" "$OUT/cfg${c}_n$N.log" | cut -c1-900; grep -E "Error|error|Traceback|Maximum resident|Elapsed" "$OUT/cfg${c}_n$N.log" | tail -5
done
