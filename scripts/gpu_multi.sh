# N-GPU check: multi tests, then the bench under torchrun at N = $1
N=${1:-2}
mkdir -p gpurun_out
python -m pytest tests/test_gpu_multi.py tests/test_gpu_kernels.py -q -m gpu -x 2>&1 | tail -5
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo rc=$?
python - <<PY
import json
d=json.load(open("gpurun_out/bench_n$N.json"))
print({k:d.get(k) for k in ("value","ms_per_step","window_ms_per_step","window_spread","gpu_launches","n_gpus")})
print(d["e2e"]); print(d["roofline"])
PY
tail -3 gpurun_out/bench_n$N.err
