"""Does the REFERENCE reproduce itself?  (TEST INFRASTRUCTURE; build container only.)

Runs the unmodified reference loop over its own StlWrapper
(oracle/_ref/librails_ref_stl.so, built by oracle/build_ref.py from
/root/reference/src) on the parity cases

  * with 1, 2, 3, 4 and 8 OpenBLAS threads (a different summation order inside
    dgemm/dgemv, nothing else), and
  * compiled with FMA contraction (-O2 -mfma -ffp-contract=fast: what any FMA
    machine, a GPU included, does to `a*b + c`; the reference's own Release flags
    are -O3 -ffast-math -march=native, CMakeLists.txt:18),

and records iteration counts, space sizes and the largest relative difference of
the residual history against the 1-thread default build.  A case whose count or
history moves under these rounding-level changes cannot be held to "same
iteration count, history to 1e-8" against ANY other implementation; the GPU
parity tests read tests/golden/ref_self_stability.json to decide which bar a case
gets, and profiles/r02_parity.md quotes it.

    python scripts/ref_self_stability.py            # writes the json
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

L10 = {"Tolerance": 1e-4, "Expand size": 1, "Lanczos iterations": 10}
CASES = {
    # the golden cases (tests/golden/make_golden.py)
    "lap2d_g16": ("laplacian", 16, 2, 1, dict(L10)),
    "C1": ("laplacian", 64, 2, 1, dict(L10)),
    "lap2d_g24_restart": ("laplacian", 24, 2, 2, {"Tolerance": 1e-5, "Expand size": 3, "Lanczos iterations": 4,
                                                  "Restart size": 40, "Reduced size": 20}),
    "convdiff2d_g20": ("convdiff", 20, 2, 3, {"Tolerance": 1e-6, "Expand size": 2, "Lanczos iterations": 5,
                                              "Restart iterations": 15}),
    "lap3d_g8": ("laplacian", 8, 3, 1, {"Tolerance": 1e-6, "Expand size": 1, "Lanczos iterations": 10,
                                        "Minimize solution space": 0}),
    # cases of tests/test_gpu_solve.py and tests/test_gpu_lanczos.py
    "lap2d_g32": ("laplacian", 32, 2, 1, dict(L10)),
    "lap2d_g40_tol1e-5": ("laplacian", 40, 2, 1, {"Tolerance": 1e-5, "Expand size": 1, "Lanczos iterations": 10}),
    "convdiff3d_g12_n3": ("convdiff", 12, 3, 3, {"Tolerance": 1e-5, "Expand size": 1, "Lanczos iterations": 10}),
    "lap2d_g128": ("laplacian", 128, 2, 1, dict(L10)),
    "lap3d_g24_restart": ("laplacian", 24, 3, 1, {"Tolerance": 1e-5, "Expand size": 1, "Lanczos iterations": 12,
                                                  "Restart size": 60, "Reduced size": 30}),
    "convdiff3d_g20": ("convdiff", 20, 3, 1, {"Tolerance": 1e-4, "Expand size": 1, "Lanczos iterations": 10,
                                              "Minimize solution space": 0}),
    "lap3d_g20_n2_L5": ("laplacian", 20, 3, 2, {"Tolerance": 1e-5, "Expand size": 1, "Lanczos iterations": 5}),
    "convdiff3d_g14_n3_e2": ("convdiff", 14, 3, 3, {"Tolerance": 1e-5, "Expand size": 2, "Lanczos iterations": 7,
                                                    "Restart size": 50, "Reduced size": 25}),
    # fixed iteration count, no restart (test_no_restart_history_and_solution_to_1e8)
    "lap2d_g48_12it": ("laplacian", 48, 2, 1, {"Tolerance": 1e-12, "Maximum iterations": 12, "Expand size": 1,
                                               "Lanczos iterations": 10, "Restart iterations": -1,
                                               "Minimize solution space": 0}),
    # rank of the residual operator (2k + n) >= L from the first iteration on: no breakdown
    "lap2d_g48_n4_L4_12it": ("laplacian", 48, 2, 4, {"Tolerance": 1e-12, "Maximum iterations": 12, "Expand size": 1,
                                                     "Lanczos iterations": 4, "Restart iterations": -1,
                                                     "Minimize solution space": 0}),
    "lap2d_g32_n2_L4": ("laplacian", 32, 2, 2, {"Tolerance": 1e-4, "Expand size": 1, "Lanczos iterations": 4}),
}
FMA_LIB = "/tmp/librails_ref_stl_fma.so"


def worker():
    import harness
    from rails_b200 import problems
    out = {}
    for name, (kind, g, dim, n_b, P) in CASES.items():
        rp, ci, v = problems.build(kind, g, dim)
        r = harness.ref_stl_solve(rp, ci, v, n_b, P, v_cap_cols=400)
        out[name] = {"ret": int(r["ret"]), "iterations": int(r["iterations"]), "k": int(r["k"]),
                     "rel": [float(x) for x in r["rel"]]}
    print(json.dumps(out))


def build_fma():
    import build_ref
    libdir, blas = build_ref.openblas_path()
    ref = build_ref.REF
    srcs = [os.path.join(ref, "src", f) for f in ("StlWrapper.cpp", "StlVector.cpp", "LapackWrapper.cpp",
                                                   "SlicotWrapper.cpp", "Timer.cpp")]
    subprocess.check_call(["g++", "-std=c++14", "-O2", "-mfma", "-ffp-contract=fast", "-fPIC", "-shared", "-fopenmp",
                           "-I" + ref, "-I" + os.path.join(ROOT, "oracle"), "-I" + os.path.join(ROOT, "rails_b200", "host"),
                           "-o", FMA_LIB, os.path.join(ROOT, "oracle", "ref_entry_stl.cpp")] + srcs +
                          [os.path.join(ROOT, "rails_b200", "host", "sb03md_shim.cpp"), blas,
                           "-Wl,--disable-new-dtags,-rpath," + libdir])


def main():
    import numpy as np
    build_fma()
    runs = {}
    for label, env in ([("threads=%d" % t, {"OPENBLAS_NUM_THREADS": str(t)}) for t in (1, 2, 3, 4, 8)] +
                       [("fma,threads=1", {"OPENBLAS_NUM_THREADS": "1", "RAILS_REF_STL_LIB": FMA_LIB})]):
        e = dict(os.environ)
        e.update(env)
        txt = subprocess.check_output([sys.executable, os.path.abspath(__file__), "--worker"], env=e,
                                      stderr=subprocess.DEVNULL)
        runs[label] = json.loads(txt.decode().strip().splitlines()[-1])
        print(label, {k: v["iterations"] for k, v in runs[label].items()}, flush=True)
    base = runs["threads=1"]
    out = {}
    for name in CASES:
        rec = {"iterations": {}, "k": {}, "max_rel_history_diff_vs_threads1": {}, "first_iteration_above_1e-8": {}}
        b = np.array(base[name]["rel"])
        for label, r in runs.items():
            rec["iterations"][label] = r[name]["iterations"]
            rec["k"][label] = r[name]["k"]
            o = np.array(r[name]["rel"])
            n = min(o.size, b.size)
            d = np.abs(o[:n] - b[:n]) / b[:n]
            rec["max_rel_history_diff_vs_threads1"][label] = float(d.max())
            rec["first_iteration_above_1e-8"][label] = int(np.argmax(d > 1e-8)) + 1 if np.any(d > 1e-8) else None
        counts = set(rec["iterations"].values())
        rec["self_reproducible"] = len(counts) == 1 and len(set(rec["k"].values())) == 1
        rec["iterations_min_max"] = [min(counts), max(counts)]
        rec["history_spread_same_count"] = max(
            [v for lab, v in rec["max_rel_history_diff_vs_threads1"].items()
             if rec["iterations"][lab] == rec["iterations"]["threads=1"]] or [0.0])
        # the 1-thread default-build run itself: what the GPU parity tests compare with
        rec["reference"] = {"ret": base[name]["ret"], "iterations": base[name]["iterations"],
                            "k": base[name]["k"], "rel": base[name]["rel"]}
        rec["problem"] = {"kind": CASES[name][0], "g": CASES[name][1], "dim": CASES[name][2],
                          "n_b": CASES[name][3], "params": CASES[name][4]}
        out[name] = rec
    path = os.path.join(ROOT, "tests", "golden", "ref_self_stability.json")
    json.dump({"cases": out, "how": "scripts/ref_self_stability.py: unmodified reference loop + StlWrapper, "
               "OpenBLAS 1/2/3/4/8 threads and an FMA-contracted build"}, open(path, "w"), indent=1)
    print("wrote", path)


if __name__ == "__main__":
    if "--worker" in sys.argv:
        worker()
    else:
        main()
