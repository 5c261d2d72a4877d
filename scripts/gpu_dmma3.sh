python -m pytest tests/test_gpu_kernels.py -q -m gpu -x -k "tensor_cores or gemm" 2>&1 | tail -2
for ring in 0 1; do
RAILS_GEMM_DMMA_RING=$ring python bench.py --kernels-only 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])['kernels']
for k in ('gemm_kxr_restart',):
    print('ring=$ring', k, {x:d[k][x] for x in ('ms','gbs','tflops','frac_of_fp64_measured')})
"
done
RAILS_GEMM_DMMA_RING=1 python -m pytest tests/test_gpu_kernels.py -q -m gpu -x -k "tensor_cores or gemm" 2>&1 | tail -2
