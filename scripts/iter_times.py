"""Per-iteration wall times and host section totals of one solve (diagnostic)."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rails_b200 import api, problems
g = int(sys.argv[1]) if len(sys.argv) > 1 else 256
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 25
api.init(0)
rp, ci, v = problems.laplacian(g, 3)
m = rp.size - 1
B = np.asfortranarray(np.random.RandomState(1).uniform(-1, 1, (m, 1)))
P = {"Tolerance": 1e-4, "Expand size": 1, "Lanczos iterations": 10, "Maximum iterations": iters}
r = api.solve(rp, ci, v, 1, P, B=B)
d = np.diff(np.concatenate([[0.0], r["seconds"]]))
print("fused=%s" % os.environ.get("RAILS_FUSED_LANCZOS", "1"))
print("ms per iteration:", " ".join("%.1f" % (1e3 * x) for x in d))
print("sections:", r["sections"])
print("solve_seconds", r["solve_seconds"])
