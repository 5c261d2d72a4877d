mkdir -p gpurun_out
timeout 100 python -m pytest tests/test_gpu_schur.py -m gpu -q -x > gpurun_out/pytest_schur.log 2>&1; echo pytest_schur rc=$?; tail -3 gpurun_out/pytest_schur.log
timeout 100 python scripts/schur_profile.py > gpurun_out/schur_profile.log 2>&1; echo profile rc=$?; tail -1 gpurun_out/schur_profile.log | cut -c1-1800
