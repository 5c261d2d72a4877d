"""Measurements of the singular-mass-matrix path on one B200:

  * the reference's ocean-model problem (tests/golden/moc_fixture.npz) through
    rails_host_solve_schur for the projection methods of matlab/test/test_MOC.m,
    beside the numpy restatement (iterations, history agreement, both residuals);
  * the device LU solve and the Schur operator apply on a larger synthetic
    problem (2-D convection-diffusion, half of the domain massless): time per
    call with CUDA events, achieved bytes/s against the algorithmic
    12 nnz(L+U) + 16 n c of a solve.

Writes gpurun_out/schur_profile.json (copied to profiles/ by hand)."""
import json
import os
import sys
import time

import numpy as np
import scipy.sparse as sp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)

from rails_b200 import api, problems  # noqa: E402
import rails_matlab  # noqa: E402
import rails_schur  # noqa: E402
import test_gpu_schur as tg  # noqa: E402
import test_schur_oracle as cpu  # noqa: E402


def solve_only(grid):
    """One small command for ncu: the factors of the large A11, three solves."""
    rp, ci, v = problems.convection_diffusion(grid, 2)
    n = rp.size - 1
    A = sp.csr_matrix((v, ci, rp))
    A11 = sp.csr_matrix(A[: n // 2][:, : n // 2])
    A11.sort_indices()
    lu = api.SparseLU(A11.indptr, A11.indices, A11.data)
    b = np.random.RandomState(0).standard_normal((A11.shape[0], int(os.environ.get("SCHUR_NRHS", "4"))))
    x = lu.device_solve(b, repeat=int(os.environ.get("SCHUR_REPEAT", "2")))
    print("solve_only", lu.device_stats, float(np.linalg.norm(A11 @ x - b) / np.linalg.norm(b)))


def main():
    args = [a for a in sys.argv[1:] if a != "solve"]
    grid = int(args[0]) if args else 256
    api.init(0)
    api.host()
    if "solve" in sys.argv[1:]:
        return solve_only(grid)
    out = {"moc": [], "follow": [], "large": {}}

    p = cpu.moc()
    A2, m2, B2 = rails_schur.add_border(p["A"], p["mdiag"], p["B"], p["border"])
    sc = rails_schur.Schur(A2, m2, B2)
    for pm, method in [(1, 0), (2.2, 0), (2.2, 1), (1.2, 1), (2.3, 0)]:
        P = tg.moc_solve_params(pm, method)
        api.solve_schur(p["rowptr"], p["colidx"], p["vals"], p["mdiag"], p["B"], P, border=p["border"])
        t0 = time.time()
        q = api.solve_schur(p["rowptr"], p["colidx"], p["vals"], p["mdiag"], p["B"], P, border=p["border"])
        wall = time.time() - t0
        rec = dict(pm=pm, method=method, ret=q["ret"], iterations=q["iterations"], k=q["k"], wall_s=wall,
                   setup_s=q["setup_seconds"], solve_s=q["solve_seconds"], stats=q["stats"])
        if "V_reduced" in q:
            rec["residual_reduced"], rec["residual_full"] = cpu.moc_residuals(p, sc, q["V_reduced"], q["T"])
        out["moc"].append(rec)
        print(rec, flush=True)

    for name, pm in [("moc", 2.2), ("moc", 1.2), ("synthetic", 2.2), ("synthetic", 1.2)]:
        pr = tg.problem(name)
        tol = 1e-3 if name == "moc" else 1e-6
        if pr["border"] is not None:
            a2, mm2, b2 = rails_schur.add_border(pr["A"], pr["mdiag"], pr["B"], pr["border"])
        else:
            a2, mm2, b2 = pr["A"], pr["mdiag"], pr["B"]
        s2 = rails_schur.Schur(a2, mm2, b2)
        t0 = time.time()
        o = rails_matlab.rails_solver(s2.S, s2.MS, s2.BS, 1000, tol,
                                      {"expand": 1, "fast_orthogonalization": False,
                                       "projection_method": pm, "Ainv": s2.Sinv})
        t_oracle = time.time() - t0
        P = tg.moc_solve_params(pm, 1)
        P["Tolerance"] = tol
        q = api.solve_schur(pr["rowptr"], pr["colidx"], pr["vals"], pr["mdiag"], pr["B"], P, border=pr["border"])
        m = min(q["iterations"], o["resvec"].size)
        err = np.abs(q["rel"][:m] - o["resvec"][:m]) / o["resvec"][:m]
        rec = dict(name=name, pm=pm, ret=q["ret"], iter_gpu=q["iterations"], iter_oracle=int(o["iter"]),
                   space_equal=bool(np.array_equal(q["space"][:m], o["space"][:m])),
                   err_first8=float(np.max(err[:8])), err_max=float(np.max(err)), err=err.tolist()[:40],
                   oracle_s=t_oracle, gpu_s=q["solve_seconds"])
        if "V_reduced" in q:
            Xq = q["V_reduced"] @ q["T"] @ q["V_reduced"].T
            Xo = o["V"] @ o["T"] @ o["V"].T
            rec["x_rel_diff"] = float(np.linalg.norm(Xq - Xo) / np.linalg.norm(Xo))
        out["follow"].append(rec)
        print({k: v for k, v in rec.items() if k != "err"}, flush=True)

    # larger operator: half of a g x g convection-diffusion domain massless
    rp, ci, v = problems.convection_diffusion(grid, 2)
    n = rp.size - 1
    rs = np.random.RandomState(0)
    mdiag = rs.uniform(0.5, 1.5, n)
    mdiag[: n // 2] = 0.0
    t0 = time.time()
    op = api.SchurOperator(rp, ci, v, mdiag)
    setup = time.time() - t0
    st = op.stats()
    large = dict(grid=grid, n=n, n1=op.n1, n2=op.n2, setup_s=setup, stats=st, apply=[], solve=[])
    A = sp.csr_matrix((v, ci, rp))
    for c in (1, 4, 16):
        X = rs.standard_normal((op.n2, c))
        Y = op.apply(op.S, X, repeat=20)
        ms = op.ms_per_apply
        # check against the definition
        i1, i2 = op.idx1, op.idx2
        import scipy.sparse.linalg as spl
        ref = A[i2][:, i2] @ X - A[i2][:, i1] @ spl.spsolve(sp.csc_matrix(A[i1][:, i1]), A[i1][:, i2] @ X).reshape(op.n1, -1)
        large["apply"].append(dict(c=c, ms=ms, rel_err=float(np.linalg.norm(Y - ref) / np.linalg.norm(ref))))
    i1 = op.idx1
    A11 = sp.csr_matrix(A[i1][:, i1])
    A11.sort_indices()
    lu = api.SparseLU(A11.indptr, A11.indices, A11.data)
    for c in (1, 4, 16):
        b = rs.standard_normal((A11.shape[0], c))
        x = lu.device_solve(b, repeat=20)
        ds = lu.device_stats
        algo = 12.0 * ds["nnz"] + 16.0 * A11.shape[0] * c
        large["solve"].append(dict(c=c, ms=ds["ms_per_solve"], levels=ds["levels"],
                                   launches=ds["launches_per_solve"], nnz=ds["nnz"],
                                   algorithmic_bytes=algo, gbs=algo / ds["ms_per_solve"] / 1e6,
                                   rel_res=float(np.linalg.norm(A11 @ x - b) / np.linalg.norm(b))))
    out["large"] = large
    print(large, flush=True)
    out["gpu_launches"] = int(api.cuda().rails_cuda_kernel_launches())
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "schur_profile.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
