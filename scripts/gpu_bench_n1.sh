mkdir -p gpurun_out
python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo rc=$?
python - <<'PY'
import json
d=json.load(open("gpurun_out/bench_n1.json"))
print({k:d[k] for k in ("value","ms_per_step","window_ms_per_step","window_spread","gpu_launches","clocks")})
print(d["e2e"]); print(d["roofline"]); print(d["cpu_baseline"])
PY
python bench.py --impl reference > gpurun_out/bench_ref_n1.json 2> gpurun_out/bench_ref_n1.err; echo rc=$?; cut -c1-1500 gpurun_out/bench_ref_n1.json
