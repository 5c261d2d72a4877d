# Round-1 closing run on one B200: the whole GPU suite, smoke, the bench line and
# the reference arm, the Schur measurements, and ncu evidence for the
# triangular-solve kernels (launch list of two solves + one full capture of the
# first -- longest -- chain kernel).  ncu only after the plain command exited 0.
mkdir -p gpurun_out
timeout 560 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_final.log 2>&1; echo pytest rc=$?; tail -4 gpurun_out/pytest_gpu_final.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final.log 2>&1; echo smoke rc=$?; tail -2 gpurun_out/smoke_final.log
timeout 400 python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo bench rc=$?; cut -c1-400 gpurun_out/bench_final.json
timeout 150 python scripts/schur_profile.py > gpurun_out/schur_profile.log 2>&1; echo profile rc=$?
export SCHUR_NRHS=1 SCHUR_REPEAT=1
S="python scripts/schur_profile.py solve"
timeout 60 $S > gpurun_out/plain_sptrsv.log 2>&1 && timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_sptrsv.csv $S > gpurun_out/ncu_sptrsv.log 2>&1
echo rc_launches=$?
timeout 120 ncu --set full --clock-control none --import-source on -k regex:sptrsv_chain -s 0 -c 1 -f -o gpurun_out/prof_sptrsv_chain $S > gpurun_out/ncu_sptrsv_chain.log 2>&1
echo rc_chain=$?
timeout 200 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_final.json 2> gpurun_out/bench_ref_final.err; echo ref rc=$?
