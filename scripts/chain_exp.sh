python -m pytest tests/test_gpu_lanczos.py tests/test_gpu_solve.py -q -m gpu 2>&1 | grep -E "^E  |FAILED|passed|failed" | head -8
for k in 2 5 11 15 24 30 40 47; do echo -n "k=$k default "; python scripts/chain_bench.py --grid 256 --k $k 2>&1 | tail -1 | cut -c1-150; done
for g in 128 256; do python scripts/profile_sections.py C3 $g 2>&1 | tail -1 | cut -c1-330; done
