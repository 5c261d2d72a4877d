# full GPU check of a build: tests, N=1 bench (both arms), ncu launch list, one --set full capture of the step kernel
mkdir -p gpurun_out
python -m pytest tests -q -m gpu -x 2>&1 | tail -5 > gpurun_out/pytest_gpu.txt; cat gpurun_out/pytest_gpu.txt
python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo rc=$?
python - <<'PY'
import json
d=json.load(open("gpurun_out/bench_n1.json"))
print({k:d[k] for k in ("value","ms_per_step","window_ms_per_step","window_spread","gpu_launches","clocks")})
print(d["e2e"]); print(d["roofline"]); print(d["cpu_baseline"])
PY
if [ "$1" != "nocu" ]; then
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches.csv python bench.py --steps 10 --warmup 3 --windows 1 --no-kernels --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1; echo ncu_rc=$?
ncu --set full --clock-control none --import-source on -k regex:lanczos_step_kernel -s 20 -c 1 -o gpurun_out/step_full python scripts/chain_bench.py --grid 256 --k 15 --reps 2 > gpurun_out/ncu_full.log 2>&1; echo ncu_full_rc=$?
ncu -i gpurun_out/step_full.ncu-rep --page raw --csv > gpurun_out/step_full_raw.csv 2>/dev/null
fi
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
