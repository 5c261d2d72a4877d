mkdir -p gpurun_out
python scripts/ncu_kernels.py > gpurun_out/ncu_kernels_plain.log 2>&1; echo plain_rc=$?
ncu --set full --clock-control none --import-source on -k regex:'spmm_kernel|dot_kernel|gemm_kernel|dot_finish' -c 16 -o gpurun_out/kernels_full python scripts/ncu_kernels.py > gpurun_out/ncu_kernels.log 2>&1; echo ncu_rc=$?
ncu -i gpurun_out/kernels_full.ncu-rep --page raw --csv > gpurun_out/kernels_full_raw.csv 2>/dev/null
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/kernels_full_raw.csv', errors='ignore')))
h=rows[0]
want=['Kernel Name','launch__grid_size','gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','dram__throughput.avg.pct_of_peak_sustained_elapsed','launch__registers_per_thread','sm__warps_active.avg.pct_of_peak_sustained_active','lts__t_sector_hit_rate.pct']
idx=[h.index(w) for w in want if w in h]
print([h[i] for i in idx]); print([rows[1][i] for i in idx])
for r in rows[2:]:
    print([r[i][:60] for i in idx])
PY
