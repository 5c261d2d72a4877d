# FP64 tensor-core kernels (csrc/gram.cu): kernel tests, per-kernel timings at the
# C3 row count, then the whole GPU suite (dot and gemm dispatch feed every solve).
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q -s -k "tensor_cores or fp64_peak or dot or gemm" > gpurun_out/pytest_dmma.log 2>&1; echo pytest_dmma rc=$?; grep -E "FP64 peak|passed|failed" gpurun_out/pytest_dmma.log | tail -4
timeout 200 python bench.py --kernels-only > gpurun_out/bench_kernels.json 2> gpurun_out/bench_kernels.err; echo kernels rc=$?
python - <<'P'
import json
d=json.loads(open('gpurun_out/bench_kernels.json').read().strip().splitlines()[-1])['kernels']
for k in ('gemm_kxr_restart','dot_kxk_projection','dot_kx1','gemm_kx1_acc','cgs2_1_vs_k'):
    print(k, {a:(round(b,3) if isinstance(b,float) else b) for a,b in d[k].items()})
P
timeout 400 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_dmma.log 2>&1; echo pytest_all rc=$?; tail -3 gpurun_out/pytest_gpu_dmma.log
