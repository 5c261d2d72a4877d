python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 scripts/e2e_trace.py C3 4 2>&1 | grep -v OMP | tail -9
python scripts/e2e_trace.py C3 3 2>&1 | tail -6
python -m pytest tests/test_gpu_multi.py tests/test_gpu_kernels.py tests/test_gpu_dropin.py -q -m gpu -x 2>&1 | tail -3
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --workload C5 --grid 128 --no-kernels > gpurun_out/c5_small.json 2> gpurun_out/c5_small.err; echo rc=$?
python -c "
import json; d=json.load(open('gpurun_out/c5_small.json')); print(d['value'], d['e2e']['time_to_tolerance_s']); print(json.dumps(d['continuation']))"
tail -3 gpurun_out/c5_small.err
