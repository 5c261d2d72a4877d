N=${1:-8}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 scripts/e2e_trace.py C3 4 2>&1 | grep -E "trace|run " | tail -8
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --workload C5 --windows 3 > gpurun_out/bench_c5_n$N.json 2> gpurun_out/bench_c5_n$N.err; echo rc=$?
python - <<PY
import json
d=json.load(open("gpurun_out/bench_c5_n$N.json"))
print({k:d.get(k) for k in ("value","ms_per_step","window_ms_per_step","gpu_launches","n_gpus")})
print(d["config"]["workload"]); print(d["e2e"]); print(d["roofline"]); print(json.dumps(d["continuation"]))
PY
tail -3 gpurun_out/bench_c5_n$N.err
