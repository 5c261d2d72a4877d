mkdir -p gpurun_out
A="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-kernels"
$A > gpurun_out/plainA.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_fused.csv $A > gpurun_out/ncuA.log 2>&1
echo rc=$?
cut -c1-300 gpurun_out/plainA.log
