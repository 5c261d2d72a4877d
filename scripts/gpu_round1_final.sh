# Round-1 closing run on one B200: GPU tests, the bench line, and ncu evidence
# for the "Residual method" = 2 path (launch list of one solve + full captures
# of its two streaming kernels).  Run under gpurun; ncu only after the plain
# command exited 0.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_final.log 2>&1; echo pytest rc=$?; tail -3 gpurun_out/pytest_gpu_final.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final.log 2>&1; echo smoke rc=$?; tail -2 gpurun_out/smoke_final.log
python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo bench rc=$?
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_final.json 2> gpurun_out/bench_ref_final.err; echo ref rc=$?
M2="python scripts/tsqr_vs_lanczos.py C3I 0 2 1"
$M2 > gpurun_out/plain_m2.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches_m2.csv $M2 > gpurun_out/ncu_m2.log 2>&1
echo rc_launches=$?
for k in dot_kernel gemm_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 60 -c 1 -f -o gpurun_out/prof_m2_$k $M2 > gpurun_out/ncu_m2_$k.log 2>&1
  echo rc_$k=$?
done
