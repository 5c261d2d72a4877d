"""Bench window (W=5 + K=20 iterations) of one workload at a given grid size:
device time per step, launches per step and the section table.  Run it once
plain and once with RAILS_PROFILE=1 in the environment (every section fenced, so
the table shows GPU time per section instead of host time).

    python scripts/profile_sections.py C3 128     # m = 128^3 = the per-GPU rows of C3 on 8 GPUs
"""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rails_b200 import api, problems
name = sys.argv[1] if len(sys.argv) > 1 else "C3"
kind, g, dim, n_b, params = problems.CONFIGS[name]
if len(sys.argv) > 2:
    g = int(sys.argv[2])
import ctypes
world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
dist = None
if world > 1:
    import torch.distributed as dist
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group(backend="gloo", rank=rank, world_size=world)
api.init(int(os.environ.get("LOCAL_RANK", "0"))); api.host()
if world > 1:
    ident = [None]
    if rank == 0:
        buf = ctypes.create_string_buffer(128)
        api.check(api.cuda().rails_cuda_comm_unique_id(buf), "comm_unique_id")
        ident[0] = buf.raw
    dist.broadcast_object_list(ident, src=0)
    api.check(api.cuda().rails_cuda_comm_init(rank, world, ident[0]), "comm_init")
m = g ** dim
r0, r1 = problems.partition_rows(m, world, rank)
rp, ci, v = problems.build(kind, g, dim, r0, r1)
M = problems.lumped_mass(m, r0, r1) if params.get("Use mass matrix") else None
B = problems.reference_B(m, n_b, r0, r1)
_solve = api.solve
api.solve = lambda *a, **k: _solve(*a, global_rows=m, row_offset=r0, **k)
W, K = 5, 20
P = dict(params); P["Tolerance"] = 0.0; P["Maximum iterations"] = W + K; P["Skip random draws"] = 1
api.solve(rp, ci, v, n_b, P, B=B, M=M)
runs = []
for _ in range(5):
    runs.append(api.solve(rp, ci, v, n_b, P, B=B, M=M, time_from=W, time_to=W + K))
runs.sort(key=lambda r: r["timed_ms"])
best = runs[len(runs) // 2]
sec = dict(kv.split("=") for kv in best["sections"].split(";") if kv)
if rank == 0:
  print(json.dumps({"rows": r1 - r0, "ranks": world, "profile": os.environ.get("RAILS_PROFILE", "0"),
                  "alpha": os.environ.get("RAILS_LANCZOS_ALPHA", "projected"),
                  "ms_per_step_median": best["timed_ms"] / K, "ms_per_step_all": [r["timed_ms"] / K for r in runs],
                  "launches_per_step": best["timed_launches"] / K, "host_syncs_per_step": best["timed_syncs"] / K,
                  "sections_ms_per_iteration": {k: round(1e3 * float(x) / (W + K), 4) for k, x in sec.items()}}))
