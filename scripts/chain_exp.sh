python -m pytest tests/test_gpu_lanczos.py tests/test_gpu_solve.py tests/test_gpu_dropin.py -q -m gpu 2>&1 | grep -E "^E  |FAILED|passed|failed" | head
for a in projected dot; do for g in 128 256; do for k in 15 50; do
  RAILS_LANCZOS_ALPHA=$a python scripts/chain_bench.py --grid $g --k $k 2>&1 | tail -1 | cut -c1-330
done; done; done
for a in projected dot; do for g in 128 256; do RAILS_LANCZOS_ALPHA=$a python scripts/profile_sections.py C3 $g 2>&1 | tail -1 | cut -c1-420; done; done
