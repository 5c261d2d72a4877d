"""Device-resident residual-Lanczos chain (rails_mv_lanczos_run) timed on its own:
L steps + one read-back, exactly as the solver issues it.

    python scripts/chain_bench.py [--rows R | --grid G] [--k K] [--n N] [--L L] [--reps 10]

Prints ms per Lanczos step and the algorithmic-bandwidth figure
16 m (2k + n) + 56 m bytes per step.  Used to size the
kernels for the per-GPU row counts of the 8-GPU runs (m / 8 = 2.1 M rows), where
a kernel lasts ~80 us and tails, launch gaps and the exchanges decide.
RAILS_LANCZOS_ALPHA=dot selects the two-exchange step for A/B.
"""
import argparse
import ctypes
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.environ.get("RAILS_PKG_ROOT", ROOT))     # A/B against another build of the package
from rails_b200 import api  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=0)
ap.add_argument("--grid", type=int, default=256)
ap.add_argument("--k", type=int, default=15)
ap.add_argument("--n", type=int, default=1)
ap.add_argument("--L", type=int, default=10)
ap.add_argument("--reps", type=int, default=10)
a = ap.parse_args()
rows = a.rows or a.grid ** 3
api.init(0)
lib = api.cuda()
P = ctypes.c_void_p
k, n, L = a.k, a.n, a.L
V = api.MultiVector(rows, k)
AV = api.MultiVector(rows, k)
B = api.MultiVector(rows, n)
Q = api.MultiVector(rows, L + 1)
for mv, c, s in ((V, k, 3), (AV, k, 4), (B, n, 5)):
    api.check(lib.rails_mv_random(mv.h, 0, c, 0, ctypes.c_uint(s)), "random")
    api.check(lib.rails_mv_scal(mv.h, 0, c, 1.0 / np.sqrt(rows)), "scal")
T = np.asfortranarray(-1e-2 * np.eye(k))
hs = (P * 3)(V.h, AV.h, B.h)
c0 = (ctypes.c_int * 3)(0, 0, 0)
nn = (ctypes.c_int * 3)(k, k, n)
al, be = np.zeros(L), np.zeros(L)
steps = ctypes.c_int(0)


def run():
    api.check(lib.rails_mv_random(Q.h, 0, 1, 0, ctypes.c_uint(9)), "random")
    api.check(lib.rails_mv_lanczos_run(3, hs, c0, nn, api.dptr(T), k, Q.h, 0, L, api.dptr(al), api.dptr(be),
                                       ctypes.byref(steps)), "lanczos_run")


for _ in range(3):
    run()
l0 = lib.rails_cuda_kernel_launches()
api.check(lib.rails_cuda_event_record(2), "event")
for _ in range(a.reps):
    run()
api.check(lib.rails_cuda_event_record(3), "event")
ms = ctypes.c_float(0)
api.check(lib.rails_cuda_event_elapsed_ms(2, 3, ctypes.byref(ms)), "elapsed")
launches = (lib.rails_cuda_kernel_launches() - l0) / a.reps
per_step = (ms.value / a.reps) / L
byts = 16.0 * rows * (2 * k + n) + 56.0 * rows
print(json.dumps({"rows": rows, "k": k, "n": n, "L": L, "ms_per_step": per_step,
                  "algorithmic_gbs": byts / per_step / 1e6, "launches_per_chain": launches,
                  "steps_done": steps.value, "alpha": os.environ.get("RAILS_LANCZOS_ALPHA", "projected"),
                  "alpha_values": [float(x) for x in al[:3]]}))
