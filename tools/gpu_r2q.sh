#!/bin/bash
# 8-GPU visit: bench at N (parity + device time line), then the full-size config 5 through
# tests/full_config.py.  Usage: bash tools/gpu_r2q.sh <tag> "<list of N>" [cfg]
TAG=${1:-r02q}; NS=${2:-"8"}; CFG=${3:-5}
OUT=gpurun_out/$TAG; mkdir -p "$OUT"
nvidia-smi -L > "$OUT/gpus.txt"; free -g | head -2 >> "$OUT/gpus.txt"
bash tools/gpu_dbg2.sh $TAG "$NS"
if [ -n "$CFG" ]; then
  N=$(echo $NS | awk '{print $NF}')
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 \
    tests/full_config.py --cfg $CFG --window-bases 200e6 --out "$OUT/full_cfg${CFG}_n$N.json" > "$OUT/full_cfg${CFG}_n$N.log" 2>&1
  echo "== cfg $CFG N=$N exit $?"; grep "^{" "$OUT/full_cfg${CFG}_n$N.log" | cut -c1-1200; grep -E "Error|error|Traceback" "$OUT/full_cfg${CFG}_n$N.log" | tail -5
fi
