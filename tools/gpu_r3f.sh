#!/bin/bash
# Last 1-GPU visit of round 3: all -m gpu tests on HEAD, the bench line (equal_work, DRAM-basis roofline),
# then the CLI end-to-end timing (tools/cli_e2e.py).  Usage: bash tools/gpu_r3f.sh <tag>
TAG=${1:-r03f}; OUT=gpurun_out/$TAG; mkdir -p "$OUT"
timeout 150 python -m pytest tests -m gpu -x -q > "$OUT/pytest_gpu.log" 2>&1; echo "pytest exit $?" | tee -a "$OUT/pytest_gpu.log"; tail -4 "$OUT/pytest_gpu.log"
timeout 120 python bench.py --steps 20 --warmup 3 > "$OUT/bench.json" 2> "$OUT/bench.err"; echo "bench exit $?"; python tools/benchsum.py < "$OUT/bench.json"; tail -3 "$OUT/bench.err"
timeout 100 python tools/cli_e2e.py > "$OUT/cli_e2e.json" 2> "$OUT/cli_e2e.err"; echo "cli_e2e exit $?"; cat "$OUT/cli_e2e.json"; tail -3 "$OUT/cli_e2e.err"
