mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_final.log 2>&1; echo pytest rc=$?; tail -3 gpurun_out/pytest_gpu_final.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final.log 2>&1; echo smoke rc=$?; tail -1 gpurun_out/smoke_final.log
