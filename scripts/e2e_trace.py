"""End-to-end one-shot solve (rails_host_solve_x, host buffers in and out) of a
workload, a few times, with the phase trace on: where the wall clock goes.

    [torchrun --nproc-per-node N ...] python scripts/e2e_trace.py [C3] [repeats]
"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rails_b200 import api, problems
name = sys.argv[1] if len(sys.argv) > 1 else "C3"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
dist = None
if world > 1:
    import torch.distributed as dist
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group(backend="gloo", rank=rank, world_size=world)
import ctypes
kind, g, dim, n_b, params = problems.CONFIGS[name]
m = g ** dim
r0, r1 = problems.partition_rows(m, world, rank)
api.init(local)
lib = api.cuda()
api.host()
if world > 1:
    ident = [None]
    if rank == 0:
        buf = ctypes.create_string_buffer(128)
        api.check(lib.rails_cuda_comm_unique_id(buf), "comm_unique_id")
        ident[0] = buf.raw
    dist.broadcast_object_list(ident, src=0)
    api.check(lib.rails_cuda_comm_init(rank, world, ident[0]), "comm_init")
rp, ci, v = problems.build(kind, g, dim, r0, r1)
M = problems.lumped_mass(m, r0, r1) if params.get("Use mass matrix") else None
B = problems.reference_B(m, n_b, r0, r1)
P = dict(params); P["Skip random draws"] = 1
os.environ["RAILS_E2E_TRACE"] = "1" if rank == 0 else ""
if rank != 0:
    del os.environ["RAILS_E2E_TRACE"]
for i in range(reps):
    if dist is not None:
        dist.barrier()
    t0 = time.time()
    out = api.solve(rp, ci, v, n_b, P, B=B, global_rows=m, row_offset=r0, v_cap_cols=int(os.environ.get("RAILS_TRACE_VCAP", "64")), M=M)
    t1 = time.time()
    if dist is not None:
        dist.barrier()
    t2 = time.time()
    if rank == 0:
        print("run %d: api.solve %.4f s (+ barrier %.4f s), ret %d, iterations %d, space %d (max %d), setup %.4f, solve %.4f" % (
            i, t1 - t0, t2 - t1, out["ret"], out["iterations"], out["k"], int(out["space"].max()), out["setup_seconds"], out["solve_seconds"]), flush=True)
