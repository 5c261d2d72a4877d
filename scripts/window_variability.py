"""Why do identical bench windows differ (round-1 driver BENCH 36.3 ms/step vs
SCALE N=1 18.6 ms/step, same kernel times)?  Runs the bench's timed window with
and without an nvidia-smi poller at several periods and prints, per window, the
device time per step, the host-side per-iteration deltas and the section totals.

    python scripts/window_variability.py [C3] [reps]
"""
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rails_b200 import api, problems  # noqa: E402
import bench  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "C3"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
api.init(0)
api.host()
lib = api.cuda()
kind, g, dim, n_b, params = problems.CONFIGS[name]
m = g ** dim
rp, ci, v = problems.build(kind, g, dim)
M = problems.lumped_mass(m) if params.get("Use mass matrix") else None
B = np.asfortranarray(np.random.RandomState(1).uniform(-1, 1, (m, n_b)))
W, K = 5, 20
P = bench.window_params(params, W, K)
Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
     "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
     "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

api.solve(rp, ci, v, n_b, P, B=B, M=M)      # cold
out = []
for mode in ("none", "smi200", "none", "smi1000", "smi200", "none"):
    proc = None
    if mode != "none":
        proc = subprocess.Popen(["nvidia-smi", "-i", "0", "--query-gpu=" + Q, "--format=csv,noheader,nounits",
                                 "-lms", mode[3:]], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        time.sleep(0.5)
    for r_ in range(reps):
        lib.rails_cuda_sync()
        t0 = time.time()
        r = api.solve(rp, ci, v, n_b, P, B=B, M=M, time_from=W, time_to=W + K)
        wall = time.time() - t0
        d = np.diff(np.concatenate([[0.0], r["seconds"]])) * 1e3
        rec = {"mode": mode, "ms_per_step": r["timed_ms"] / K, "solve_s": r["solve_seconds"],
               "setup_s": r["setup_seconds"], "wall_s": wall,
               "iter_ms_median": float(np.median(d[W:])), "iter_ms_max": float(np.max(d[W:])),
               "iter_ms": [round(float(x), 2) for x in d], "sections": r["sections"]}
        out.append(rec)
        print(mode, "%.2f ms/step" % rec["ms_per_step"], "host median %.2f max %.2f" %
              (rec["iter_ms_median"], rec["iter_ms_max"]), flush=True)
    if proc:
        proc.terminate()
        proc.wait()
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "window_variability.json"), "w"), indent=1)
