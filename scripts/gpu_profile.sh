mkdir -p gpurun_out
set -x
A="python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-kernels"
B="python bench.py --kernels-only --kernel-reps 2"
$A > gpurun_out/plainA.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_r01.csv $A > gpurun_out/ncuA.log 2>&1
echo rcA=$?
$B > gpurun_out/plainB.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:dot_kernel -s 3 -c 1 -f -o gpurun_out/prof_dot $B > gpurun_out/ncuB1.log 2>&1
echo rcB1=$?
ncu --set full --clock-control none --import-source on -k regex:gemm_kernel -s 3 -c 1 -f -o gpurun_out/prof_gemm $B > gpurun_out/ncuB2.log 2>&1
echo rcB2=$?
ncu --set full --clock-control none --import-source on -k regex:spmm_kernel -s 3 -c 1 -f -o gpurun_out/prof_spmm $B > gpurun_out/ncuB3.log 2>&1
echo rcB3=$?
tail -3 gpurun_out/ncuA.log gpurun_out/ncuB1.log
ls -la gpurun_out
