mkdir -p gpurun_out/r01d
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r01d/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -25 gpurun_out/r01d/pytest_gpu.log
echo "== bench N=1"
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu 2>&1 | python tools/benchsum.py
echo "== bench N=2"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 10 --warmup 3 2>&1 | grep -v "^W1\|^\*\*\*" | python tools/benchsum.py
echo "== bench N=2 mirror=0"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --steps 10 --warmup 3 --opt mirror=0 2>&1 | grep -v "^W1\|^\*\*\*" | python tools/benchsum.py
