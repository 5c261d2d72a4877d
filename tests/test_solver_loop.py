"""Pins the solver loop the product ships (rails_b200/host/RailsSolver.hpp) to
the reference loop (/root/reference/src/LyapunovSolver.hpp): both instantiated
with the reference's own StlWrapper for all three concepts must produce the
same bits.  Needs oracle/_ref (built where /root/reference exists)."""
import numpy as np
import pytest

import harness
from rails_b200 import problems

pytestmark = pytest.mark.skipif(harness.ref() is None, reason="oracle/_ref not built")

CASES = [
    ("laplacian", 16, 2, 1, {"Tolerance": 1e-4, "Expand size": 1, "Lanczos iterations": 10}),
    ("laplacian", 12, 2, 2, {"Tolerance": 1e-6, "Expand size": 2, "Lanczos iterations": 9}),
    ("convdiff", 14, 2, 3, {"Tolerance": 1e-5, "Expand size": 3, "Lanczos iterations": 8,
                            "Restart size": 30, "Reduced size": 12}),
    ("laplacian", 6, 3, 1, {"Tolerance": 1e-6, "Expand size": 1, "Lanczos iterations": 6,
                            "Restart iterations": 7, "Minimize solution space": 0}),
    ("laplacian", 5, 2, 1, {"Tolerance": 1e-12, "Expand size": 1, "Lanczos iterations": 4,
                            "Maximum iterations": 9}),
]


@pytest.mark.parametrize("kind,g,dim,n_b,P", CASES)
def test_product_loop_is_bit_identical_to_reference_loop(kind, g, dim, n_b, P):
    rp, ci, v = problems.build(kind, g, dim)
    r = harness.ref_stl_solve(rp, ci, v, n_b, P, v_cap_cols=200)
    p = harness.ref_stl_solve(rp, ci, v, n_b, P, v_cap_cols=200, entry="b200loop_stl_solve")
    assert (p["ret"], p["iterations"], p["k"]) == (r["ret"], r["iterations"], r["k"])
    assert np.array_equal(p["abs"], r["abs"])
    assert np.array_equal(p["rel"], r["rel"])
    assert np.array_equal(p["V"], r["V"])
    assert np.array_equal(p["T"], r["T"])


def test_parameter_rejection_matches():
    rp, ci, v = problems.laplacian(6, 2)
    P = {"Expand size": 4, "Lanczos iterations": 4}
    r = harness.ref_stl_solve(rp, ci, v, 1, P)
    p = harness.ref_stl_solve(rp, ci, v, 1, P, entry="b200loop_stl_solve")
    assert r["ret"] == p["ret"] == 101


@pytest.mark.parametrize("kind,g,dim,n_b,stop", [("laplacian", 10, 2, 1, 4), ("laplacian", 10, 2, 2, 7),
                                                 ("convdiff", 6, 3, 3, 5)])
def test_tsqr_residual_estimate_is_the_exact_residual_norm(kind, g, dim, n_b, stop):
    """a17 ("Residual method" = 1): the residual's eigenpairs from a tall-skinny
    QR of [V, AV, B] and a small eigenproblem.  Neither reference implementation
    has this, so it is validated against the dense residual: after
    `stop` iterations the reported estimate must be the spectral norm of
    A X + X A' + B B' with X = V T V' (product loop over the reference's own
    StlWrapper, CPU)."""
    rp, ci, v = problems.build(kind, g, dim)
    m = rp.size - 1
    B = np.asfortranarray(np.random.RandomState(3).uniform(-1, 1, (m, n_b)))
    P = {"Tolerance": 1e-14, "Expand size": 1, "Lanczos iterations": 8, "Maximum iterations": stop,
         "Residual method": 1, "Minimize solution space": 0}
    p = harness.ref_stl_solve(rp, ci, v, n_b, P, B=B, v_cap_cols=200, entry="b200loop_stl_solve")
    assert p["ret"] == -1 and p["iterations"] == stop
    A = harness.csr_to_dense(rp, ci, v, m)
    X = p["V"] @ p["T"] @ p["V"].T
    R = A @ X + X @ A.T + B @ B.T
    exact = np.max(np.abs(np.linalg.eigvalsh(0.5 * (R + R.T))))
    assert abs(p["abs"][-1] - exact) <= 1e-9 * exact, (p["abs"][-1], exact)
    # the estimates decrease towards convergence and the method converges
    P2 = dict(P); P2["Tolerance"] = 1e-8; P2["Maximum iterations"] = 200
    q = harness.ref_stl_solve(rp, ci, v, n_b, P2, B=B, v_cap_cols=200, entry="b200loop_stl_solve")
    assert q["ret"] == 0
    X = q["V"] @ q["T"] @ q["V"].T
    R = A @ X + X @ A.T + B @ B.T
    assert np.max(np.abs(np.linalg.eigvalsh(0.5 * (R + R.T)))) <= 1e-8 * np.linalg.norm(B.T @ B, 2) * 1.01
