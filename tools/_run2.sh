mkdir -p gpurun_out/r01d
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r01d/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/r01d/pytest_gpu.log
for s in 64 32 16 8; do
echo "== bench N=1 stride $s"
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --opt anchor_stride=$s 2>&1 | python tools/benchsum.py
done
for s in 64 32 16; do
echo "== bench N=2 stride $s"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 295$s bench.py --gpus 2 --steps 10 --warmup 3 --opt anchor_stride=$s 2>&1 | grep "^{" | python tools/benchsum.py
done
