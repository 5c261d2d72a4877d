"""Time to tolerance of one solve with the reference's Lanczos estimate
("Residual method" = 0), with the incremental tall-skinny QR of [V, AV, B]
("Residual method" = 1) and with the whole iteration in the coordinates of that
basis ("Residual method" = 2) on the same inputs (diagnostic, one GPU).

    python scripts/tsqr_vs_lanczos.py [workload=C3] [grid] [methods=0,1,2] [repeats=2]"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rails_b200 import api, problems

name = sys.argv[1] if len(sys.argv) > 1 else "C3"
kind, g, dim, n_b, P = problems.CONFIGS[name]
if len(sys.argv) > 2 and int(sys.argv[2]) > 0:
    g = int(sys.argv[2])
methods = [int(x) for x in sys.argv[3].split(",")] if len(sys.argv) > 3 else [0, 1, 2]
repeats = int(sys.argv[4]) if len(sys.argv) > 4 else 2
api.init(0)
rp, ci, v = problems.build(kind, g, dim)
m = rp.size - 1
B = np.asfortranarray(np.random.RandomState(1).uniform(-1, 1, (m, n_b)))
M = problems.lumped_mass(m) if P.get("Use mass matrix") else None
out = {"workload": "%s g=%d m=%d n_b=%d" % (name, g, m, n_b)}
for method in methods * repeats:
    Q = dict(P)
    Q["Residual method"] = method
    r = api.solve(rp, ci, v, n_b, Q, B=B, M=M, v_cap_cols=0)
    d = np.diff(np.concatenate([[0.0], r["seconds"]]))
    out.setdefault("method%d" % method, []).append({
        "ret": r["ret"], "iterations": r["iterations"], "k": r["k"], "solve_seconds": r["solve_seconds"],
        "final_rel": float(r["rel"][-1]), "sections": r["sections"],
        "space": [int(x) for x in r["space"]],
        "ms_per_iteration": [round(1e3 * x, 2) for x in d]})
print(json.dumps(out))
