mkdir -p gpurun_out
for w in "$@"; do
  case $w in
    C4) extra="--parity-cpu-grid 40";;
    *) extra="";;
  esac
  timeout 1500 python bench.py --parity-full --workload $w $extra > gpurun_out/parity_full_$w.json 2> gpurun_out/parity_full_$w.err; echo "$w rc=$?"
  cut -c1-1800 gpurun_out/parity_full_$w.json; tail -2 gpurun_out/parity_full_$w.err
done
