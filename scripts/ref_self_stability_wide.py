"""Does the REFERENCE reproduce itself on the parameter lists of BASELINE configs 2 and 4?
(TEST INFRASTRUCTURE; build container only -- needs oracle/_ref/librails_ref_csr.so, the
unmodified reference loop over a CSR operator, built by oracle/build_ref.py.)

Runs that loop on reduced grids with config 2's list (n = 4, Expand size 4, 20 Lanczos
steps, restart 200/100) and config 4's (n = 16, Expand size 16, 40 Lanczos steps) with
1, 2, 4 and 8 BLAS/OpenMP threads -- a different summation order, nothing else -- and
records iteration counts, final space sizes, the relative difference of the residual
histories and of X = V T V' against the 1-thread run, plus the CPU port (oracle) with 4
threads.  Output: profiles/r02_ref_self_stability_wide.json (quoted by
profiles/r02_parity.md).

    python scripts/ref_self_stability_wide.py
"""
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

CASES = {
    "C2_list_lap2d_g64": ("laplacian", 64, 2, 4, {"Tolerance": 1e-4, "Restart size": 200, "Reduced size": 100,
                                                  "Expand size": 4, "Lanczos iterations": 20,
                                                  "Restart iterations": -1}),
    "C4_list_convdiff3d_g16": ("convdiff", 16, 3, 16, {"Tolerance": 1e-4, "Expand size": 16,
                                                       "Lanczos iterations": 40}),
}


def worker(which, case, out_prefix):
    import harness
    from rails_b200 import problems
    kind, g, dim, n_b, P = CASES[case]
    P = dict(P)
    P["Skip random draws"] = 1
    rp, ci, v = problems.build(kind, g, dim)
    B = problems.reference_B(g ** dim, n_b)
    if which == "ref":
        o = harness.ref_csr_solve(rp, ci, v, n_b, P, B=B, v_cap_cols=420)
    else:
        harness.oracle().oracle_set_num_threads(int(os.environ.get("OMP_NUM_THREADS", "1")))
        o = harness.oracle_solve(rp, ci, v, n_b, P, B=B, v_cap_cols=420)
    np.save(out_prefix + "_V.npy", o["V"])
    np.save(out_prefix + "_T.npy", o["T"])
    print(json.dumps({"iterations": int(o["iterations"]), "k": int(o["k"]), "rel": [float(x) for x in o["rel"]]}))


def main():
    import harness
    out = {}
    for case in CASES:
        runs = {}
        for which, threads in [("ref", 1), ("ref", 2), ("ref", 4), ("ref", 8), ("port", 4)]:
            label = "%s,threads=%d" % ("reference loop" if which == "ref" else "cpu port", threads)
            prefix = "/tmp/rss_wide_%s_%s_%d" % (case, which, threads)
            env = dict(os.environ, OMP_NUM_THREADS=str(threads), OPENBLAS_NUM_THREADS=str(threads))
            txt = subprocess.check_output([sys.executable, os.path.abspath(__file__), "--worker", which, case, prefix],
                                          env=env, stderr=subprocess.DEVNULL)
            r = json.loads(txt.decode().strip().splitlines()[-1])
            r["V"], r["T"] = np.load(prefix + "_V.npy"), np.load(prefix + "_T.npy")
            runs[label] = r
            print(case, label, r["iterations"], r["k"], flush=True)
        base = runs["reference loop,threads=1"]
        rec = {"problem": {"kind": CASES[case][0], "g": CASES[case][1], "dim": CASES[case][2], "n_b": CASES[case][3],
                           "params": CASES[case][4]}, "runs": {}}
        b = np.array(base["rel"])
        for label, r in runs.items():
            o = np.array(r["rel"])
            n = min(o.size, b.size)
            d = np.abs(o[:n] - b[:n]) / b[:n]
            rec["runs"][label] = {
                "iterations": r["iterations"], "k": r["k"],
                "history_rel_diff_first6": [float(x) for x in d[:6]],
                "history_rel_diff_max": float(d.max()),
                "first_iteration_above_1e-8": int(np.argmax(d > 1e-8)) + 1 if np.any(d > 1e-8) else None,
                "x_rel_diff_vs_reference_1_thread": float(harness.lowrank_rel_diff(r["V"], r["T"], base["V"], base["T"]))}
        out[case] = rec
    path = os.path.join(ROOT, "profiles", "r02_ref_self_stability_wide.json")
    json.dump({"how": "scripts/ref_self_stability_wide.py: unmodified reference loop over a CSR operator, "
                      "1/2/4/8 BLAS+OpenMP threads; CPU port with 4", "cases": out}, open(path, "w"), indent=1)
    print("wrote", path)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "--worker":
        worker(sys.argv[2], sys.argv[3], sys.argv[4])
    else:
        main()
