"""Measured parity of the GPU solve against the REFERENCE ITSELF, case by case.

For every case of tests/golden/ref_self_stability.json (the unmodified reference
loop + StlWrapper, 1 BLAS thread; its own spread over thread counts and an
FMA-contracted build is recorded beside it) the solve runs on the GPU through
rails_host_solve in the three Lanczos modes

    chain   device-resident recurrence (default)
    step    fused kernels, k x k products on the host
    plain   operation by operation, the reference's call sequence

and the script records iteration count, space size, the largest relative
difference of the residual history (first 10 iterations / all) and of the
solution.  Output: gpurun_out/parity_report.json and a markdown table on stdout
(committed as profiles/r02_parity.md).  No /root/reference needed at run time.
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from rails_b200 import api, problems  # noqa: E402

MODES = (("chain", "1"), ("step", "step"), ("plain", "0"))


def main():
    api.init(0)
    api.host()
    stab = json.load(open(os.path.join(ROOT, "tests", "golden", "ref_self_stability.json")))["cases"]
    golden = np.load(os.path.join(ROOT, "tests", "golden", "ref_golden.npz"))
    out = {}
    for name, rec in stab.items():
        pr, ref = rec["problem"], rec["reference"]
        rp, ci, v = problems.build(pr["kind"], pr["g"], pr["dim"])
        ref_rel = np.array(ref["rel"])
        row = {"reference": {"iterations": ref["iterations"], "k": ref["k"],
                             "own_iterations_min_max": rec["iterations_min_max"],
                             "own_history_spread": rec["history_spread_same_count"],
                             "self_reproducible": rec["self_reproducible"]}}
        for mode, env in MODES:
            os.environ["RAILS_FUSED_LANCZOS"] = env
            r = api.solve(rp, ci, v, pr["n_b"], pr["params"], v_cap_cols=400)
            n = min(r["iterations"], ref["iterations"])
            d = np.abs(r["rel"][:n] - ref_rel[:n]) / ref_rel[:n]
            m = {"ret": int(r["ret"]), "iterations": int(r["iterations"]), "k": int(r["k"]),
                 "history_diff_first10": float(d[:10].max()), "history_diff_all": float(d.max()),
                 "first_iteration_above_1e-8": int(np.argmax(d > 1e-8)) + 1 if np.any(d > 1e-8) else None}
            if name + "_Xcorner" in golden and "V" in r:
                Xs = (r["V"][:64] @ r["T"]) @ r["V"][:64].T
                g = golden[name + "_Xcorner"]
                m["x_corner_diff"] = float(np.linalg.norm(Xs - g) / np.linalg.norm(g))
            row[mode] = m
        del os.environ["RAILS_FUSED_LANCZOS"]
        out[name] = row
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "parity_report.json"), "w"), indent=1)
    print("| case | reference: iterations (own range), k, own history spread | mode | iterations, k | "
          "history diff, first 10 | history diff, all | first iteration > 1e-8 | X corner diff |")
    print("|---|---|---|---|---|---|---|---|")
    for name, row in out.items():
        rf = row["reference"]
        for mode, _ in MODES:
            m = row[mode]
            print("| %s | %d (%d-%d), %d, %.1e | %s | %d, %d | %.1e | %.1e | %s | %s |" % (
                name, rf["iterations"], rf["own_iterations_min_max"][0], rf["own_iterations_min_max"][1], rf["k"],
                rf["own_history_spread"], mode, m["iterations"], m["k"], m["history_diff_first10"],
                m["history_diff_all"], m["first_iteration_above_1e-8"],
                "%.1e" % m["x_corner_diff"] if "x_corner_diff" in m else "-"))


if __name__ == "__main__":
    main()
