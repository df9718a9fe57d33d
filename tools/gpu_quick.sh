#!/bin/bash
# Quick GPU visit: parity tests + bench (no ncu).  Usage: bash tools/gpu_quick.sh <tag> [bench args...]
TAG=${1:-q}; shift
OUT=gpurun_out/$TAG; mkdir -p "$OUT"
timeout 900 python -m pytest tests -m gpu -x -q > "$OUT/pytest_gpu.log" 2>&1; echo "pytest exit $?" >> "$OUT/pytest_gpu.log"; tail -4 "$OUT/pytest_gpu.log"
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu "$@" > "$OUT/bench.json" 2> "$OUT/bench.err"; echo "bench exit $?"
python tools/benchsum.py < "$OUT/bench.json"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > "$OUT/smoke.log" 2>&1; echo "smoke exit $?"; tail -2 "$OUT/smoke.log"; tail -3 "$OUT/bench.err"
