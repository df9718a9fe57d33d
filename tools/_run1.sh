mkdir -p gpurun_out/r01c
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r01c/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/r01c/pytest_gpu.log
for o in "anchor_stride=64" "anchor_stride=32" "anchor_stride=16"; do
  echo "== $o"
  timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu --opt $o 2>&1 | python tools/benchsum.py
done
