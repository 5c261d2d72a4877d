python -m pytest tests/test_gpu_kernels.py -q -m gpu -x -k "transpose or halo_plan or round_trip" 2>&1 | tail -15
timeout 300 python bench.py --parity-full --workload C2 --grid 256 > gpurun_out/parity_full_C2_g256.json 2> gpurun_out/parity_full_C2_g256.err; echo rc=$?; cut -c1-1500 gpurun_out/parity_full_C2_g256.json
python scripts/profile_sections.py C3 2>&1 | tail -1 | cut -c1-400
