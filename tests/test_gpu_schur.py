"""Singular mass matrix on the GPU: the device LU solves
(rails_lu_*, csrc/sptrsv.cu), the row gather/scatter, the Schur-complement
operator (rails_b200/host/CudaSchurOperator.cpp) and the solve through it
(rails_host_solve_schur), all through the C ABI / the host entry points.

Checkers: scipy's SuperLU and A x = b itself for the solves;
oracle/rails_schur.py (restatement of matlab/RAILSschur.m) for the operator
pieces; oracle/rails_matlab.py with the projection methods for the solve; and
the reference's own bar for this path, matlab/test/test_MOC.m:16-90, on the
reference's ocean-model fixture."""
import ctypes
import os
import sys

import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spl

import test_schur_oracle as cpu
from rails_b200 import problems

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import rails_matlab  # noqa: E402
import rails_schur  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["random40", "random300", "random1500", "laplace2d", "convdiff3d",
                                  "saddle", "moc_A11", "moc_A"])
@pytest.mark.parametrize("nrhs", [1, 3, 7])
def test_device_lu_solve(gpu, name, nrhs):
    A = cpu.lu_cases()[name]
    A.sort_indices()
    n = A.shape[0]
    lu = gpu.SparseLU(A.indptr, A.indices, A.data)
    assert lu.info == 0
    b = np.random.RandomState(nrhs).uniform(-1, 1, (n, nrhs))
    for trans in (False, True):
        xh = lu.solve(b, trans)                   # same factors, host arithmetic
        xd = lu.device_solve(b, trans)
        # same factors, same operations up to the summation order inside a row
        scale = np.linalg.norm(xh, axis=0)
        assert np.max(np.linalg.norm(xd - xh, axis=0) / scale) < 1e-9, name
        Aop = A.T if trans else A
        res = np.linalg.norm(Aop @ xd - b) / (spl.norm(A, 1) * np.linalg.norm(xd) + np.linalg.norm(b))
        resh = np.linalg.norm(Aop @ xh - b) / (spl.norm(A, 1) * np.linalg.norm(xh) + np.linalg.norm(b))
        assert res < max(10 * resh, 1e-14)
        st = lu.device_stats
        assert st["nnz"] == lu.nnz_l + lu.nnz_u
        # the level chains collapse into a handful of launches
        assert st["launches_per_solve"] <= 2 + 2 * (2 * max(st["levels"]) // 3 + 2)
        # the schedule the kernels walk, executed on the host, is the same solve
        xe = lu.emulate_device_schedule(b, trans)
        assert np.max(np.linalg.norm(xd - xe, axis=0) / scale) < 1e-9, name
    # bit-reproducible
    assert np.array_equal(lu.device_solve(b), lu.device_solve(b))


def test_device_lu_levels_are_a_few_launches(gpu):
    """A 2-D Laplacian under a fill-reducing ordering: hundreds of levels, yet a solve
    is a few dozen kernels (runs of small levels are walked by one CTA)."""
    rp, ci, v = problems.laplacian(96, 2)
    lu = gpu.SparseLU(rp, ci, v)
    b = np.random.RandomState(0).uniform(-1, 1, (rp.size - 1, 4))
    x = lu.device_solve(b, repeat=5)
    A = sp.csr_matrix((v, ci, rp))
    assert np.linalg.norm(A @ x - b) < 1e-10 * np.linalg.norm(b)
    st = lu.device_stats
    assert max(st["levels"]) > 200
    assert st["launches_per_solve"] < max(st["levels"]) // 3


def test_gather_scatter_rows(gpu):
    lib = gpu.cuda()
    rs = np.random.RandomState(5)
    n, c = 1000, 3
    idx = np.sort(rs.choice(n, 377, replace=False)).astype(np.int32)
    X = rs.standard_normal((n, c))
    I = ctypes.c_void_p()
    gpu.check(lib.rails_idx_create(idx.size, gpu.iptr(idx), ctypes.byref(I)), "idx_create")
    src = gpu.MultiVector.from_numpy(X)
    dst = gpu.MultiVector(idx.size, c)
    gpu.check(lib.rails_mv_gather_rows(dst.h, 0, src.h, 0, c, I, 1.0), "gather")
    assert np.array_equal(dst.numpy(), X[idx])
    back = gpu.MultiVector.from_numpy(np.full((n, c), 7.0))
    gpu.check(lib.rails_mv_scatter_rows(back.h, 0, dst.h, 0, c, I, -2.0), "scatter")
    want = np.full((n, c), 7.0)
    want[idx] = -2.0 * X[idx]
    assert np.array_equal(back.numpy(), want)
    # shape errors are reported, not executed
    assert lib.rails_mv_gather_rows(dst.h, 0, dst.h, 0, c, I, 1.0) != 0
    lib.rails_idx_free(I)


def synthetic_singular(g=24, dim=2, share=0.4, seed=2):
    """Convection-diffusion operator with a random 40 % of the unknowns massless."""
    rp, ci, v = problems.convection_diffusion(g, dim)
    n = rp.size - 1
    rs = np.random.RandomState(seed)
    mdiag = rs.uniform(0.5, 1.5, n)
    massless = rs.rand(n) < share
    massless[0] = True
    mdiag[massless] = 0.0
    B = rs.uniform(-1, 1, (n, 2))
    B[massless] = 0.0
    return dict(rowptr=rp, colidx=ci, vals=v, A=sp.csr_matrix((v, ci, rp)), mdiag=mdiag, B=B, n=n,
                border=None)


def problem(name):
    if name == "moc":
        return cpu.moc()
    return synthetic_singular()


@pytest.mark.parametrize("name", ["moc", "synthetic"])
def test_schur_operator_pieces_follow_restatement(gpu, name):
    p = problem(name)
    if p["border"] is not None:
        A2, m2, B2 = rails_schur.add_border(p["A"], p["mdiag"], p["B"], p["border"])
    else:
        A2, m2, B2 = p["A"], p["mdiag"], p["B"]
    sc = rails_schur.Schur(A2, m2, B2)
    op = gpu.SchurOperator(p["rowptr"], p["colidx"], p["vals"], p["mdiag"], border=p["border"])
    assert (op.n_total, op.n1, op.n2) == (A2.shape[0], sc.idx1.size, sc.idx2.size)
    assert np.array_equal(op.idx1, sc.idx1) and np.array_equal(op.idx2, sc.idx2)
    rs = np.random.RandomState(7)
    X = rs.standard_normal((op.n2, 5))

    def close(got, want, tol):
        return np.linalg.norm(got - want) <= tol * np.linalg.norm(want)

    assert close(op.apply(op.S, X), sc.S(X), 1e-9)
    assert close(op.apply(op.ST, X), sc.S(X, True), 1e-9)
    assert close(op.apply(op.PROLONGATE, X), sc.prolongate(X), 1e-9)
    Y = op.apply(op.SINV, X)
    assert close(Y, sc.Sinv(X), 1e-7)
    assert close(op.apply(op.S, Y), X, 1e-7)                       # S S^-1 = I
    # restrict: B as given (zero in the singular part: plain selection), and a B
    # that is not (the A21 A11^-1 correction, RAILSschur.m:70-72)
    assert np.array_equal(op.apply(op.RESTRICT, B2), B2[sc.idx2])
    F = rs.standard_normal((op.n_total, 2))
    assert close(op.apply(op.RESTRICT, F), sc.restrict(F), 1e-9)
    st = op.stats()
    assert st["mvps"] == 5 + 5 + 5 and st["factor_nnz"] > 0


def moc_solve_params(pm, method=1):
    P = {"Tolerance": 1e-3, "Maximum iterations": 1000, "Expand size": 1, "Lanczos iterations": 10,
         "Restart iterations": -1, "Relative restart tolerance": 1, "Orthogonalize candidates": 1,
         "Residual method": method}
    if pm != 1:
        P["Projection method"] = pm
    return P


@pytest.mark.parametrize("pm,method", [(1, 0), (2.2, 0), (2.2, 1), (1.2, 1)])
def test_moc_solve_meets_the_reference_bar(gpu, pm, method):
    """matlab/test/test_MOC.m:16-90 through rails_host_solve_schur: the residual
    of the reduced equation and of the ORIGINAL equation (with the prolongated
    basis) are both below 1e-3."""
    p = cpu.moc()
    q = gpu.solve_schur(p["rowptr"], p["colidx"], p["vals"], p["mdiag"], p["B"], moc_solve_params(pm, method),
                        border=p["border"])
    assert q["ret"] == 0, q["ret"]
    assert (q["n1"], q["n2"]) == (1026, 512)
    A2, m2, B2 = rails_schur.add_border(p["A"], p["mdiag"], p["B"], p["border"])
    sc = rails_schur.Schur(A2, m2, B2)
    V, T = q["V_reduced"], q["T"]
    reduced, full = cpu.moc_residuals(p, sc, V, T)
    assert reduced < 1e-3 and full < 1e-3
    # the prolongation was done on the device
    n = p["n"]
    Vf = sc.Vtrans(V)
    assert np.linalg.norm(q["V"] - Vf) < 1e-8 * np.linalg.norm(Vf)
    X = q["V"][:n] @ T @ q["V"][:n].T
    A, M, B = p["A"], sp.diags(p["mdiag"]), p["B"]
    assert np.linalg.norm(A @ X @ M.T + M @ (A @ X).T + B @ B.T, "fro") < 1e-3
    if pm > 2:
        # the extended space pays: test_MOC_inv needs a third of the iterations
        assert q["iterations"] < 80


@pytest.mark.parametrize("name,pm", [("moc", 2.2), ("moc", 1.2), ("synthetic", 2.2), ("synthetic", 1.2)])
def test_schur_solve_follows_matlab_restatement(gpu, name, pm):
    """Exact residual eigenpairs ("Residual method" = 1) and a start from B: neither
    side draws a random number, so product and restatement walk the same path."""
    p = problem(name)
    tol = 1e-3 if name == "moc" else 1e-6
    if p["border"] is not None:
        A2, m2, B2 = rails_schur.add_border(p["A"], p["mdiag"], p["B"], p["border"])
    else:
        A2, m2, B2 = p["A"], p["mdiag"], p["B"]
    sc = rails_schur.Schur(A2, m2, B2)
    opts = {"expand": 1, "fast_orthogonalization": False}
    P = moc_solve_params(pm, 1)
    P["Tolerance"] = tol
    opts.update(projection_method=pm, Ainv=sc.Sinv)
    o = rails_matlab.rails_solver(sc.S, sc.MS, sc.BS, 1000, tol, opts)
    q = gpu.solve_schur(p["rowptr"], p["colidx"], p["vals"], p["mdiag"], p["B"], P, border=p["border"])
    assert q["ret"] == 0 and o["res"] < tol
    assert abs(q["iterations"] - o["iter"]) <= 1
    m = min(8, q["iterations"], o["resvec"].size)
    assert np.array_equal(q["space"][:m], o["space"][:m])
    err = np.abs(q["rel"][:m] - o["resvec"][:m]) / o["resvec"][:m]
    assert np.max(err) < 1e-5, err
    Xq = q["V_reduced"] @ q["T"] @ q["V_reduced"].T
    Xo = o["V"] @ o["T"] @ o["V"].T
    assert np.linalg.norm(Xq - Xo) / np.linalg.norm(Xo) < 5e-2      # both within tol of the solution


def test_schur_warm_start_and_driver_continuation(gpu, tmp_path):
    """Continuation as in test/LyapunovSolver_test.cpp:312-352, through the Schur
    complement: a solve started from the (prolongated) solution of the same
    problem converges at once -- through the entry point and through V.mtx."""
    import re
    import subprocess
    p = synthetic_singular(g=16)
    P = {"Tolerance": 1e-6, "Expand size": 2, "Lanczos iterations": 10, "Maximum iterations": 300}
    first = gpu.solve_schur(p["rowptr"], p["colidx"], p["vals"], p["mdiag"], p["B"], P)
    assert first["ret"] == 0 and first["iterations"] > 5
    again = gpu.solve_schur(p["rowptr"], p["colidx"], p["vals"], p["mdiag"], p["B"],
                            dict(P, **{"Restart from solution": 1}), V0=first["V"])
    assert again["ret"] == 0 and again["iterations"] <= 5
    missing = gpu.solve_schur(p["rowptr"], p["colidx"], p["vals"], p["mdiag"], p["B"],
                              dict(P, **{"Restart from solution": 1}))
    assert missing["ret"] == 102

    d = str(tmp_path)
    n = p["n"]
    gpu.mm_write_crs(os.path.join(d, "A.mtx"), p["rowptr"], p["colidx"], p["vals"], n)
    Bs = sp.csr_matrix(p["B"])
    gpu.mm_write_crs(os.path.join(d, "B.mtx"), Bs.indptr.astype(np.int32), Bs.indices.astype(np.int32),
                     Bs.data, p["B"].shape[1])
    Ms = sp.csr_matrix(sp.diags(p["mdiag"]))
    Ms.eliminate_zeros()
    gpu.mm_write_crs(os.path.join(d, "M.mtx"), Ms.indptr.astype(np.int32), Ms.indices.astype(np.int32),
                     Ms.data, n)
    ini = "[Lyapunov Solver]\nTolerance = 1e-6\nExpand size = 2\nLanczos iterations = 10\n"
    open(os.path.join(d, "p1.ini"), "w").write(ini)
    open(os.path.join(d, "p2.ini"), "w").write(ini + "Restart from solution = true\n")
    exe = os.path.join(ROOT, "rails_b200", "rails_cuda_main")
    r1 = subprocess.run([exe, os.path.join(d, "p1.ini"), "--dir", d], stdout=subprocess.PIPE,
                        stderr=subprocess.PIPE, text=True, timeout=600)
    assert r1.returncode == 0, r1.stdout[-2000:] + r1.stderr[-2000:]
    it1 = int(re.search(r"Iterations: (\d+)", r1.stdout).group(1))
    r2 = subprocess.run([exe, os.path.join(d, "p2.ini"), "--dir", d], stdout=subprocess.PIPE,
                        stderr=subprocess.PIPE, text=True, timeout=600)
    assert r2.returncode == 0, r2.stdout[-2000:] + r2.stderr[-2000:]
    it2 = int(re.search(r"Iterations: (\d+)", r2.stdout).group(1))
    assert it1 > 5 and it2 <= 5
    V = gpu.mm_read_dense(os.path.join(d, "V.mtx"))
    assert V.shape[0] == n


def test_schur_refuses_what_it_cannot_do(gpu):
    p = synthetic_singular(g=8)
    # singular A11: a massless unknown with an empty row/column block
    vals = p["vals"].copy()
    A = sp.csr_matrix((vals, p["colidx"], p["rowptr"])).tolil()
    A[0, :] = 0.0
    A[:, 0] = 0.0
    A = sp.csr_matrix(A)
    A.sort_indices()
    with pytest.raises(gpu.RailsError):
        gpu.SchurOperator(A.indptr, A.indices, A.data, p["mdiag"])
    q = gpu.solve_schur(p["rowptr"], p["colidx"], p["vals"], p["mdiag"], p["B"],
                        {"Tolerance": 1e-4, "Projection method": 3.5})
    assert q["ret"] == 101


MOC_XML = """<ParameterList name="RAILS">
  <ParameterList name="Lyapunov Solver">
    <Parameter name="Tolerance" type="double" value="1e-3"/>
    <Parameter name="Expand size" type="int" value="1"/>
    <Parameter name="Lanczos iterations" type="int" value="10"/>
    <Parameter name="Restart iterations" type="int" value="-1"/>
    <Parameter name="Projection method" type="double" value="2.2"/>
  </ParameterList>
  <ParameterList name="Eigenvalue Solver">
    <Parameter name="Number of Eigenvalues" type="int" value="5"/>
  </ParameterList>
</ParameterList>
"""


def test_cli_reduces_a_singular_mass_matrix(gpu, tmp_path):
    """rails_cuda_main on the ocean-model problem, as the reference's main.cpp runs
    it (src/main.cpp:75-116): A.mtx, B.mtx, M.mtx with zero rows in M -> Schur
    complement -> solve -> V.mtx (prolongated), T.mtx, eigenvalues of the full X."""
    import subprocess
    d = str(tmp_path)
    p = cpu.moc()
    A2, m2, B2 = rails_schur.add_border(p["A"], p["mdiag"], p["B"], p["border"])
    A2 = sp.csr_matrix(A2)
    A2.sort_indices()
    n = A2.shape[0]
    gpu.mm_write_crs(os.path.join(d, "A.mtx"), A2.indptr.astype(np.int32), A2.indices.astype(np.int32),
                     A2.data, n)
    Bs = sp.csr_matrix(B2)
    gpu.mm_write_crs(os.path.join(d, "B.mtx"), Bs.indptr.astype(np.int32), Bs.indices.astype(np.int32),
                     Bs.data, B2.shape[1])
    Ms = sp.csr_matrix(sp.diags(m2))
    Ms.eliminate_zeros()
    gpu.mm_write_crs(os.path.join(d, "M.mtx"), Ms.indptr.astype(np.int32), Ms.indices.astype(np.int32),
                     Ms.data, n)
    open(os.path.join(d, "params.xml"), "w").write(MOC_XML)
    exe = os.path.join(ROOT, "rails_b200", "rails_cuda_main")
    r = subprocess.run([exe, os.path.join(d, "params.xml"), "--dir", d], stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "Computing Schur complement" in r.stdout
    assert "Singular part: 1026 rows, Schur complement: 512 rows" in r.stdout
    assert "Amount of matrix-vector products after the solve" in r.stdout
    V = gpu.mm_read_dense(os.path.join(d, "V.mtx"))
    T = gpu.mm_read_dense(os.path.join(d, "T.mtx"))
    assert V.shape[0] == n and T.shape == (V.shape[1], V.shape[1])
    # residual of the ORIGINAL equation (test_MOC.m:28-33)
    no = p["n"]
    X = V[:no] @ T @ V[:no].T
    A, M, B = p["A"], sp.diags(p["mdiag"]), p["B"]
    assert np.linalg.norm(A @ X @ M.T + M @ (A @ X).T + B @ B.T, "fro") < 1e-3
    # printed eigenvalues: those of the full X = V T V' (V is not orthonormal here)
    Xf = V @ T @ V.T
    lam = np.linalg.eigvalsh(0.5 * (Xf + Xf.T))
    lam = lam[np.argsort(-np.abs(lam))][:5]
    lines = r.stdout.split("Eigenvalues of X")[1].splitlines()[1:6]
    shown = np.array([[float(x) for x in line.split()] for line in lines])
    assert np.allclose(shown[:, 0], lam, rtol=1e-4)
    assert np.allclose(shown[:, 1], lam / np.trace(Xf), rtol=1e-4)
