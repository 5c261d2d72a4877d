N=${1:-8}
for p in 0 1; do RAILS_PROFILE=$p python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$p scripts/profile_sections.py C3 2>&1 | grep '^{' | tail -1; done
nproc
