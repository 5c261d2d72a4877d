mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_multi.py -m gpu -q > gpurun_out/pytest_multi2.log 2>&1; echo pytest_multi rc=$?; tail -3 gpurun_out/pytest_multi2.log
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err; echo bench2 rc=$?; tail -1 gpurun_out/bench_2gpu.json | cut -c1-260
