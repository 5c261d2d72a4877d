mkdir -p gpurun_out
for rep in 1 2; do
for mode in 1 0; do
  RAILS_FP64_TENSOR=$mode timeout 200 python bench.py --no-cpu-baseline --no-kernels > gpurun_out/ab_${mode}_${rep}.json 2> gpurun_out/ab_${mode}_${rep}.err
  python - <<P
import json
d=json.loads(open('gpurun_out/ab_${mode}_${rep}.json').read().strip().splitlines()[-1])
print('tensor=$mode rep=$rep value %.2f ms/step %.2f e2e solve %.3f tsqr solve %.3f launches %d' % (d['value'], d['ms_per_step'], d['e2e']['solve_seconds'], d['e2e_tsqr']['solve_seconds'], d['gpu_launches']))
P
done
done
