# After the FP64 tensor-core kernels: kernel tests again (gemm pipeline changed),
# kernel timings, the bench line, the restart-heavy C2 workload, ncu captures of
# the two new kernels.
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "tensor_cores or gemm" > gpurun_out/pytest_dmma.log 2>&1; echo pytest_dmma rc=$?; tail -1 gpurun_out/pytest_dmma.log
timeout 200 python bench.py --kernels-only > gpurun_out/bench_kernels.json 2> gpurun_out/bench_kernels.err; echo kernels rc=$?
python - <<'P'
import json
d=json.loads(open('gpurun_out/bench_kernels.json').read().strip().splitlines()[-1])['kernels']
for k in ('gemm_kxr_restart','dot_kxk_projection'):
    print(k, {a:(round(b,3) if isinstance(b,float) else b) for a,b in d[k].items()})
P
timeout 400 python bench.py > gpurun_out/bench_final2.json 2> gpurun_out/bench_final2.err; echo bench rc=$?; cut -c1-200 gpurun_out/bench_final2.json
timeout 200 python bench.py --workload C2 --no-cpu-baseline --no-kernels > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo c2 rc=$?; cut -c1-300 gpurun_out/bench_c2.json
K="python bench.py --kernels-only --kernel-reps 2"
for k in gram_dmma gemm_dmma; do
  timeout 200 ncu --set full --clock-control none --import-source on -k regex:$k -s 2 -c 1 -f -o gpurun_out/prof_$k $K > gpurun_out/ncu_$k.log 2>&1
  echo rc_$k=$?
done
