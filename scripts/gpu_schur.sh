# N4 on one B200: the Schur / LU GPU tests and the command-line driver tests, the
# Schur measurements, then ncu evidence for the triangular-solve kernels (launch
# list of three solves + one full capture of the chain kernel; ncu only after the
# plain command exited 0).
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_schur.py tests/test_gpu_cli.py -m gpu -q > gpurun_out/pytest_schur.log 2>&1; echo pytest_schur rc=$?; tail -25 gpurun_out/pytest_schur.log
timeout 420 python scripts/schur_profile.py > gpurun_out/schur_profile.log 2>&1; echo profile rc=$?; tail -4 gpurun_out/schur_profile.log | cut -c1-1500
S="python scripts/schur_profile.py solve"
timeout 120 $S > gpurun_out/plain_sptrsv.log 2>&1 && timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_sptrsv.csv $S > gpurun_out/ncu_sptrsv.log 2>&1
echo rc_launches=$?
timeout 300 ncu --set full --clock-control none --import-source on -k regex:sptrsv_chain -s 40 -c 1 -f -o gpurun_out/prof_sptrsv_chain $S > gpurun_out/ncu_sptrsv_chain.log 2>&1
echo rc_chain=$?
