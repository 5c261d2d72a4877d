"""Singular mass matrix, CPU side:

  * oracle/rails_schur.py + the projection methods of oracle/rails_matlab.py are
    pinned to the reference's own tests of this path, matlab/test/test_MOC.m:
    16-90, on the reference's ocean-model fixture (tests/golden/moc_fixture.npz,
    converted by tests/golden/make_moc_fixture.py): both explicit residuals
    below 1e-3 for the plain solve and for projection method 2.2 with Sinv;
  * the product's host factorisation (rails_b200/host/SparseLU.cpp, the stand-in
    for Amesos_Klu) against scipy's SuperLU and against A x = b itself;
  * the projection methods in the product loop (over the reference's StlWrapper,
    CPU) walk the same path as the restatement.

The device side is tests/test_gpu_schur.py."""
import os
import sys

import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spl

import harness
from rails_b200 import api, problems

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import rails_matlab  # noqa: E402
import rails_schur  # noqa: E402


def moc():
    d = np.load(os.path.join(ROOT, "tests", "golden", "moc_fixture.npz"))
    A = sp.csr_matrix((d["vals"], d["colidx"], d["rowptr"]))
    return dict(A=A, rowptr=d["rowptr"], colidx=d["colidx"], vals=d["vals"], mdiag=d["mdiag"],
                B=d["B"], border=d["border"], n=A.shape[0])


def moc_residuals(p, sc, V, T):
    """The two assertions of test_MOC.m:25-33."""
    n = p["n"]
    SV, MS = sc.S(V), sc.MS
    reduced = np.linalg.norm(SV @ T @ (V.T @ MS.T) + (MS @ V) @ (SV @ T).T + sc.BS @ sc.BS.T, "fro")
    Vf = sc.Vtrans(V)[:n]
    A, M, B = p["A"], sp.diags(p["mdiag"]), p["B"]
    X = Vf @ T @ Vf.T
    full = np.linalg.norm(A @ X @ M.T + M @ (A @ X).T + B @ B.T, "fro")
    return reduced, full


def test_fixture_is_the_reference_problem():
    p = moc()
    assert p["n"] == 8 * 8 * 4 * 6                               # test_MOC.m:19
    assert p["A"].nnz == 17364
    assert int(np.sum(p["mdiag"] == 0.0)) == 4 * p["n"] // 6     # u, v, w, p carry no mass
    assert np.count_nonzero(p["B"]) > 0 and not np.any(p["B"].reshape(-1, 6)[:, :5])


@pytest.mark.parametrize("opts", [{}, {"projection_method": 2.2}])
def test_restatement_meets_the_reference_inequalities(opts):
    p = moc()
    A2, m2, B2 = rails_schur.add_border(p["A"], p["mdiag"], p["B"], p["border"])
    sc = rails_schur.Schur(A2, m2, B2)
    assert sc.idx1.size == 1026 and sc.idx2.size == 512 and not sc.B_in_singular_part
    opts = dict(opts)
    if "projection_method" in opts:
        opts["Ainv"] = sc.Sinv                                   # test_MOC.m:45-49
    np.random.seed(1)
    r = rails_matlab.rails_solver(sc.S, sc.MS, sc.BS, 1000, 1e-3, opts)
    assert r["res"] < 1e-3
    reduced, full = moc_residuals(p, sc, r["V"], r["T"])
    assert reduced < 1e-3 and full < 1e-3                        # test_MOC.m:26,33


def test_schur_restatement_identities():
    """S S^-1 = I, S' is the transpose of S, the prolongation satisfies the
    singular rows of the original system."""
    p = moc()
    A2, m2, B2 = rails_schur.add_border(p["A"], p["mdiag"], p["B"], p["border"])
    sc = rails_schur.Schur(A2, m2, B2)
    rs = np.random.RandomState(0)
    x, y = rs.standard_normal((sc.idx2.size, 3)), rs.standard_normal((sc.idx2.size, 2))
    assert np.linalg.norm(sc.S(sc.Sinv(x)) - x) < 1e-8 * np.linalg.norm(x)
    assert abs(np.sum(y[:, :2] * sc.S(x[:, :2])) - np.sum(sc.S(y[:, :2], True) * x[:, :2])) < 1e-10
    full = sc.prolongate(x)
    assert np.linalg.norm((A2 @ full)[sc.idx1]) < 1e-9 * np.linalg.norm(full)
    assert np.linalg.norm((A2 @ full)[sc.idx2] - sc.S(x)) < 1e-9 * np.linalg.norm(x)


# ---- the host factorisation ---------------------------------------------------
def lu_cases():
    rs = np.random.RandomState(3)
    out = {}
    for n, dens in [(40, 0.15), (300, 0.02), (1500, 0.003)]:
        A = sp.random(n, n, dens, random_state=rs, format="csr") + sp.diags(rs.uniform(-1, 1, n))
        out["random%d" % n] = sp.csr_matrix(A)
    rp, ci, v = problems.laplacian(40, 2)
    out["laplace2d"] = sp.csr_matrix((v, ci, rp))
    rp, ci, v = problems.convection_diffusion(10, 3)
    out["convdiff3d"] = sp.csr_matrix((v, ci, rp))
    # saddle point: zero diagonal block forces off-diagonal pivots
    K = sp.csr_matrix((v, ci, rp))
    C = sp.random(60, K.shape[0], 0.02, random_state=rs, format="csr") + sp.eye(60, K.shape[0])
    out["saddle"] = sp.csr_matrix(sp.bmat([[K, C.T], [C, None]]))
    p = moc()
    A2, m2, _ = rails_schur.add_border(p["A"], p["mdiag"], p["B"], p["border"])
    i1 = np.where(m2 == 0.0)[0]
    out["moc_A11"] = sp.csr_matrix(A2[i1][:, i1])
    out["moc_A"] = sp.csr_matrix(A2)
    return out


@pytest.mark.parametrize("name", ["random40", "random300", "random1500", "laplace2d", "convdiff3d",
                                  "saddle", "moc_A11", "moc_A"])
def test_host_lu_solves(name):
    A = lu_cases()[name]
    A.sort_indices()
    n = A.shape[0]
    lu = api.SparseLU(A.indptr, A.indices, A.data)
    assert lu.info == 0
    pinv, q = lu.permutations()
    assert sorted(pinv.tolist()) == list(range(n)) and sorted(q.tolist()) == list(range(n))
    rs = np.random.RandomState(1)
    b = rs.uniform(-1, 1, (n, 3))
    ref = spl.splu(sp.csc_matrix(A))
    for trans in (False, True):
        x = lu.solve(b, trans)
        Aop = A.T if trans else A
        xr = ref.solve(b, trans="T" if trans else "N")
        # backward error of our solve no worse than a few times SuperLU's
        ours = np.linalg.norm(Aop @ x - b) / (spl.norm(A, 1) * np.linalg.norm(x) + np.linalg.norm(b))
        theirs = np.linalg.norm(Aop @ xr - b) / (spl.norm(A, 1) * np.linalg.norm(xr) + np.linalg.norm(b))
        assert ours < max(50 * theirs, 1e-14), (ours, theirs)
        # the device schedule (chains, shared-memory slots, windows) walked on the
        # host gives the same solve: checks the plan of csrc/sptrsv.cu without a GPU
        xe = lu.emulate_device_schedule(b, trans)
        assert np.linalg.norm(xe - x) < 1e-10 * np.linalg.norm(x)
        assert all(s["levels"] >= 1 for s in lu.plan_stats)
    if name == "saddle":
        assert lu.off_diagonal_pivots > 0
    if name == "laplace2d":
        # the ordering keeps the fill well below a banded factorisation's 2 n g
        assert lu.nnz_l + lu.nnz_u < 0.5 * 2 * n * 40


def test_host_lu_reports_singular():
    A = sp.csr_matrix(np.array([[1.0, 2.0, 0.0], [2.0, 4.0, 0.0], [0.0, 0.0, 1.0]]))
    lu = api.SparseLU(A.indptr, A.indices, A.data)
    assert lu.info > 0


def test_orderings_and_their_selection():
    """Nested dissection fills less than minimum degree on a grid and more on an
    irregular graph; the factorisation takes whichever fills less (symbolic count)."""
    rp, ci, v = problems.convection_diffusion(48, 2)
    n = rp.size - 1
    for order in (api.minimum_degree(rp, ci), api.nested_dissection(rp, ci)):
        assert sorted(order.tolist()) == list(range(n))
    md = api.SparseLU(rp, ci, v, order=api.minimum_degree(rp, ci))
    nd = api.SparseLU(rp, ci, v, order=api.nested_dissection(rp, ci))
    auto = api.SparseLU(rp, ci, v)
    assert auto.nnz_l + auto.nnz_u == min(md.nnz_l + md.nnz_u, nd.nnz_l + nd.nnz_u)
    A = lu_cases()["random1500"]
    A.sort_indices()
    md = api.SparseLU(A.indptr, A.indices, A.data, order=api.minimum_degree(A.indptr, A.indices))
    nd = api.SparseLU(A.indptr, A.indices, A.data, order=api.nested_dissection(A.indptr, A.indices))
    auto = api.SparseLU(A.indptr, A.indices, A.data)
    assert md.nnz_l + md.nnz_u < nd.nnz_l + nd.nnz_u
    # pivoting moves a few entries; the choice follows the symbolic count
    assert auto.nnz_l + auto.nnz_u <= 1.05 * (md.nnz_l + md.nnz_u)
    # a disconnected graph and a tiny one go through the same code
    B = sp.block_diag([sp.csr_matrix((v, ci, rp)), sp.eye(5), sp.csr_matrix((v, ci, rp))], format="csr")
    o = api.nested_dissection(B.indptr, B.indices, leaf=50)
    assert sorted(o.tolist()) == list(range(B.shape[0]))
    lu = api.SparseLU(B.indptr, B.indices, B.data, order=o)
    b = np.ones((B.shape[0], 1))
    assert np.linalg.norm(B @ lu.solve(b) - b) < 1e-10 * np.linalg.norm(b)


def test_minimum_degree_defers_dense_rows():
    n = 400
    rp, ci, v = problems.laplacian(20, 2)
    A = sp.lil_matrix(sp.csr_matrix((v, ci, rp)))
    A[n - 7, :] = 1.0                     # a dense row/column (the border of the MOC problem)
    A[:, n - 7] = 1.0
    A = sp.csr_matrix(A)
    order = api.minimum_degree(A.indptr, A.indices)
    assert sorted(order.tolist()) == list(range(n))
    assert order[-1] == n - 7


# ---- projection methods in the product loop (CPU, StlWrapper) ----------------------
needs_ref = pytest.mark.skipif(harness.ref() is None, reason="oracle/_ref not built")


@needs_ref
@pytest.mark.parametrize("pm,expand", [(1.2, 1), (2.2, 1), (2.2, 2), (1.3, 1), (2.3, 1), (2.1, 1), (1.1, 2)])
def test_projection_methods_follow_matlab_restatement(pm, expand):
    g, n_b, tol = 12, 2, 1e-6
    n = g * g
    rp, ci, v = problems.convection_diffusion(g, 2)
    A = harness.csr_to_dense(rp, ci, v, n)
    rs = np.random.RandomState(4)
    B = rs.rand(n, n_b)
    V0 = rs.standard_normal((n, 1))
    V0 /= np.linalg.norm(V0)
    Ainv = np.linalg.inv(A)
    o = rails_matlab.rails_solver(A, None, B, maxit=300, tol=tol, opts={
        "expand": expand, "projection_method": pm, "Ainv": lambda x: Ainv @ x, "space": V0,
        "fast_orthogonalization": False})
    P = {"Tolerance": tol, "Maximum iterations": 300, "Residual method": 1, "Restart from solution": 1,
         "Restart iterations": -1, "Lanczos iterations": 10, "Orthogonalize candidates": 1,
         "Relative restart tolerance": 1, "Expand size": expand, "Projection method": pm}
    q = harness.stl_solve_ext(rp, ci, v, B, P, V0=V0, inverse=Ainv)
    assert q["ret"] == 0 and o["res"] < tol
    assert q["iterations"] == o["iter"]
    keep = np.ones(q["iterations"], dtype=bool)
    keep[1:] = q["restarted"][:-1] == 0
    assert np.array_equal(q["space"][keep], o["space"])
    err = np.abs(q["rel"][keep] - o["resvec"]) / o["resvec"]
    assert np.max(err[:8]) < 1e-6 and np.max(err) < 1e-2, err
    Xq, Xo = q["V"] @ q["T"] @ q["V"].T, o["V"] @ o["T"] @ o["V"].T
    assert np.linalg.norm(Xq - Xo) / np.linalg.norm(Xo) < 1e-5
    # the point of the inverse: far fewer iterations than the plain projection
    if pm in (1.2, 2.2) and expand == 1:
        plain = rails_matlab.rails_solver(A, None, B, maxit=300, tol=tol, opts={
            "expand": expand, "space": V0, "fast_orthogonalization": False})
        assert o["iter"] < plain["iter"]


@needs_ref
def test_projection_method_needs_an_inverse_and_a_valid_value():
    rp, ci, v = problems.laplacian(6, 2)
    B = np.ones((36, 1))
    base = {"Tolerance": 1e-4, "Expand size": 1, "Lanczos iterations": 5}
    assert harness.stl_solve_ext(rp, ci, v, B, dict(base, **{"Projection method": 2.2}))["ret"] == 1
    assert harness.stl_solve_ext(rp, ci, v, B, dict(base, **{"Projection method": 3.1}))["ret"] == 101


def test_host_lu_on_random_small_matrices():
    """Property test of the factorisation with pivoting: for random small sparse
    matrices (zero diagonals, duplicate-free, some singular) either the solve
    reproduces numpy's to the conditioning of the matrix, or the matrix is
    reported singular exactly when it is structurally or numerically rank
    deficient."""
    rs = np.random.RandomState(12345)
    solved = singular = 0
    for trial in range(300):
        n = rs.randint(1, 40)
        dens = rs.uniform(0.05, 0.5)
        D = rs.uniform(-1, 1, (n, n)) * (rs.rand(n, n) < dens)
        if trial % 3:
            D[np.arange(n), np.arange(n)] += rs.uniform(0.5, 2.0, n) * rs.choice([-1, 1], n)
        if trial % 7 == 0 and n > 2:
            D[rs.randint(n)] = 0.0                          # an empty row: singular
        A = sp.csr_matrix(D)
        A.sort_indices()
        lu = api.SparseLU(A.indptr, A.indices, A.data)
        sv = np.linalg.svd(D, compute_uv=False)
        cond = sv[0] / sv[-1] if sv[-1] > 0 else np.inf
        if lu.info != 0:
            singular += 1
            assert cond > 1e12, (trial, n, cond)
            continue
        if cond > 1e10:
            continue                                        # factorised, but nothing to compare against
        solved += 1
        b = rs.uniform(-1, 1, (n, 2))
        for trans in (False, True):
            x = lu.solve(b, trans)
            want = np.linalg.solve(D.T if trans else D, b)
            assert np.linalg.norm(x - want) <= 1e-13 * cond * np.linalg.norm(want) + 1e-14, (trial, n, cond)
            xe = lu.emulate_device_schedule(b, trans)
            assert np.linalg.norm(xe - x) <= 1e-13 * cond * np.linalg.norm(x) + 1e-14
    assert solved > 100 and singular > 10
