"""Config C5 of BASELINE.json: 3-D Laplacian 512^3 (m = 134 217 728) row-partitioned
over 8 GPUs, two-step continuation (pattern of test/LyapunovSolver_test.cpp:312-352):
solve with A, then solve with A - 0.01 I warm-started from the first solution
("Restart from solution").  Launch with torch.distributed.run, one rank per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29700 scripts/c5_continuation.py [grid]

Prints one JSON line on rank 0."""
import ctypes
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch.distributed as dist
    from rails_b200 import api, problems
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        dist.init_process_group("gloo", rank=rank, world_size=world)
    api.init(local)
    lib = api.cuda()
    api.host()
    if world > 1:
        ident = [None]
        if rank == 0:
            buf = ctypes.create_string_buffer(128)
            api.check(lib.rails_cuda_comm_unique_id(buf), "unique_id")
            ident[0] = buf.raw
        dist.broadcast_object_list(ident, src=0)
        api.check(lib.rails_cuda_comm_init(rank, world, ident[0]), "comm_init")
    kind, g, dim, n_b, params = problems.CONFIGS["C5"]
    if len(sys.argv) > 1:
        g = int(sys.argv[1])
    m = g ** dim
    r0, r1 = problems.partition_rows(m, world, rank)
    rp, ci, v = problems.build(kind, g, dim, r0, r1)
    # B: the same global random vector whatever the partition (generated in slabs)
    rs = np.random.RandomState(1)
    B = None
    chunk = 1 << 24
    pos = 0
    while pos < m:
        blk = rs.uniform(-1.0, 1.0, min(chunk, m - pos))
        lo, hi = max(pos, r0), min(pos + blk.size, r1)
        if lo < hi:
            if B is None:
                B = np.zeros((r1 - r0, n_b), order="F")
            B[lo - r0:hi - r0, 0] = blk[lo - pos:hi - pos]
        pos += blk.size
    cap = 120

    def barrier():
        lib.rails_cuda_sync()
        if world > 1:
            dist.barrier()

    barrier()
    t0 = time.time()
    first = api.solve(rp, ci, v, n_b, params, B=B, global_rows=m, row_offset=r0, v_cap_cols=cap)
    barrier()
    t1 = time.time()
    v2 = problems.shift_diagonal(rp, ci, v, -0.01, r0)
    cold = api.solve(rp, ci, v2, n_b, params, B=B, global_rows=m, row_offset=r0, v_cap_cols=cap)
    barrier()
    t2 = time.time()
    P2 = dict(params)
    P2["Restart from solution"] = 1
    warm = api.solve(rp, ci, v2, n_b, P2, B=B, global_rows=m, row_offset=r0, v_cap_cols=cap, V0=first["V"])
    barrier()
    t3 = time.time()
    if rank == 0:
        print(json.dumps({
            "config": "C5 laplacian 3D g=%d (m=%d) over %d GPU(s), continuation A -> A - 0.01 I" % (g, m, world),
            "first": {"ret": first["ret"], "iterations": first["iterations"], "k": first["k"],
                      "wall_s": t1 - t0, "solve_s": first["solve_seconds"], "rel": float(first["rel"][-1])},
            "cold": {"ret": cold["ret"], "iterations": cold["iterations"], "k": cold["k"],
                     "wall_s": t2 - t1, "solve_s": cold["solve_seconds"], "rel": float(cold["rel"][-1])},
            "warm": {"ret": warm["ret"], "iterations": warm["iterations"], "k": warm["k"],
                     "wall_s": t3 - t2, "solve_s": warm["solve_seconds"], "rel": float(warm["rel"][-1])},
            "small_reductions": "peer mailboxes" if world > 1 and lib.rails_cuda_comm_p2p() else "local/NCCL",
        }), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
