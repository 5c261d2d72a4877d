# ncu evidence for profiles/: (A) launch list of the bench command's timed solve,
# (B) one --set full capture each of the dominant kernels.  Run under gpurun
# (one GPU), only after the plain command exited 0.
mkdir -p gpurun_out
A="python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-kernels"
B="python bench.py --kernels-only --kernel-reps 2"
$A > gpurun_out/plainA.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches_r01c.csv $A > gpurun_out/ncuA.log 2>&1
echo rcA=$?
$B > gpurun_out/plainB.log 2>&1 || exit 1
for k in lanczos_dot_kernel multi_gemv_kernel spmm_kernel finish_exchange_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 3 -c 1 -f -o gpurun_out/prof_$k $B > gpurun_out/ncu_$k.log 2>&1
  echo rc_$k=$?
done
ls -la gpurun_out | tail -12
