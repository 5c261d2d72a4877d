"""Per-iteration wall-clock deltas of the bench's window solve, run twice in one
process (first solve pays one-off allocations)."""
import sys, os, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rails_b200 import api, problems
import bench
api.init(0); api.host()
kind, g, dim, n_b, params = problems.CONFIGS[sys.argv[1] if len(sys.argv) > 1 else "C3"]
m = g ** dim
rp, ci, v = problems.build(kind, g, dim)
M = problems.lumped_mass(m) if params.get("Use mass matrix") else None
B = np.asfortranarray(np.random.RandomState(1).uniform(-1, 1, (m, n_b)))
P = bench.window_params(params, 5, 20)
for rep in range(2):
    r = api.solve(rp, ci, v, n_b, P, B=B, M=M, time_from=5, time_to=25)
    d = np.diff(np.concatenate([[0.0], r["seconds"]])) * 1e3
    print("rep", rep, "timed_ms %.1f" % r["timed_ms"], "solve_s %.3f" % r["solve_seconds"])
    print(" ".join("%.1f" % x for x in d))
