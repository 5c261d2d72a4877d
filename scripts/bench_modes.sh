# A/B of the Lanczos modes on N GPUs: device-resident chain vs per-step fused
N=${1:-2}
for mode in 1 step; do
  for rep in 1 2; do
    RAILS_FUSED_LANCZOS=$mode timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2961$rep bench.py --gpus $N --steps 20 --warmup 5 --no-kernels --workload ${2:-C3} > gpurun_out/ab_${N}_${mode}_$rep.json 2> gpurun_out/ab_${N}_${mode}_$rep.err
    python - <<PY
import json
try:
    d=json.load(open("gpurun_out/ab_${N}_${mode}_$rep.json"))
    print("N=$N mode=$mode rep=$rep value %.2f ms/step %.3f e2e %.2f iters %d" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["iterations_to_tolerance"]))
except Exception as e:
    print("N=$N mode=$mode failed", e); print(open("gpurun_out/ab_${N}_${mode}_$rep.err").read()[-1500:])
PY
  done
done
