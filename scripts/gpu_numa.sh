N=${1:-8}
lscpu | grep -E "NUMA|Socket|^CPU\(s\)" | head -8
nvidia-smi topo -m 2>/dev/null | head -14
for b in 0 1; do echo "RAILS_NUMA_BIND=$b"; RAILS_NUMA_BIND=$b RAILS_NUMA_VERBOSE=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$b scripts/e2e_trace.py C3 4 2>&1 | grep -E "trace|local cpus" | tail -5; done
