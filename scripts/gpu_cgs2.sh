python -m pytest tests/test_gpu_solve.py tests/test_gpu_lanczos.py -q -m gpu 2>&1 | tail -4
python scripts/parity_report.py > gpurun_out/parity_table.md 2> gpurun_out/parity_report.err; echo rc=$?; grep -E "lap2d_g40|convdiff2d_g20 " gpurun_out/parity_table.md
