"""GPU run of tests/test_matlab_features.py: the MATLAB-only options of the
solver loop and the incremental tall-skinny QR of
[V, AV, B] ("Residual method" = 1) over the CUDA wrappers, through
rails_host_solve_x, against oracle/rails_matlab.py (numpy restatement of
matlab/RAILSsolver.m).  Same bar as on the CPU: same iteration count and space
sizes, residual history equal to rounding, same X."""
import numpy as np
import pytest

import harness
import test_matlab_features as cpu
from rails_b200 import problems

pytestmark = pytest.mark.gpu


def diag_csr(M):
    n = M.shape[0]
    return (np.arange(n + 1, dtype=np.int32), np.arange(n, dtype=np.int32), np.ascontiguousarray(np.diag(M)))


@pytest.mark.parametrize("name,method", [(c, 1) for c in cpu.CASE_NAMES] + [(c, 2) for c in cpu.METHOD2_CASES])
def test_cuda_loop_follows_matlab_restatement(gpu, name, method):
    p = cpu.setup()
    mopts, popts, uses_m, V0, N = cpu.cases(p)[name]
    if N is not None:
        cpu.project_out(p, N)
    o = cpu.run_oracle(p, mopts, uses_m, V0, N)
    P = dict(cpu.BASE)
    P.update(popts)
    P["Residual method"] = method
    q = gpu.solve(*p["csr"], p["B"].shape[1], P, B=p["B"], v_cap_cols=p["n"], V0=V0,
                  M=diag_csr(p["M"]) if uses_m else None, nullspace=N)
    assert q["ret"] == (0 if o["res"] < cpu.TOL else -1)
    # rails_host_solve reports the history without the restart flag: rebuild it
    # from the space sizes (every other iteration grows the space)
    restarted = np.zeros(q["iterations"], dtype=np.int32)
    restarted[:-1] = np.diff(q["space"]) <= 0
    q["restarted"] = restarted
    cpu.compare(o, q, p["n"])
    V = q["V"]
    if N is not None:
        assert np.linalg.norm(N.T @ V) < 1e-10
    if "M-orthogonalization" in popts:
        assert np.linalg.norm(V.T @ p["M"] @ V - np.eye(V.shape[1])) < 1e-8


@pytest.mark.parametrize("method", [1, 2])
def test_exact_residual_eigenpairs_at_every_iteration(gpu, method):
    p = cpu.setup(g=12, n_b=2)
    A, B = p["A"], p["B"]
    for stop in (2, 9, 23, 36, 40):
        P = dict(cpu.BASE, **{"Tolerance": 1e-14, "Maximum iterations": stop, "Expand size": 1,
                              "Lanczos iterations": 3, "Minimize solution space": 0,
                              "Residual method": method})
        q = gpu.solve(*p["csr"], 2, P, B=B, v_cap_cols=p["n"], V0=p["V0"])
        assert q["iterations"] == stop and q["k"] == stop
        X = q["V"] @ q["T"] @ q["V"].T
        R = A @ X + X @ A.T + B @ B.T
        exact = np.max(np.abs(np.linalg.eigvalsh(0.5 * (R + R.T))))
        assert abs(q["abs"][-1] - exact) < 1e-9 * exact, (stop, q["abs"][-1], exact)


@pytest.mark.parametrize("method", [1, 2])
@pytest.mark.parametrize("kind,g,dim,n_b,mass", [("laplacian", 48, 2, 1, False), ("laplacian", 14, 3, 2, True),
                                                 ("convdiff", 12, 3, 3, False)])
def test_tsqr_solve_converges_to_the_dense_solution(gpu, kind, g, dim, n_b, mass, method):
    """Whole solves with "Residual method" = 1 at sizes where the dense answer is
    cheap: the converged X meets the explicit-residual bar
    (matlab/test/test_Laplace.m:31-59) and the reported estimate is the true
    residual norm at the last iteration."""
    rp, ci, v = problems.build(kind, g, dim)
    m = rp.size - 1
    B = np.asfortranarray(np.random.RandomState(5).uniform(-1, 1, (m, n_b)))
    Mc = problems.lumped_mass(m) if mass else None
    P = {"Tolerance": 1e-6, "Expand size": n_b, "Lanczos iterations": 10, "Residual method": method,
         "Restart iterations": -1, "Maximum iterations": 400, "Minimize solution space": 0}
    if mass:
        P["Use mass matrix"] = 1
    q = gpu.solve(rp, ci, v, n_b, P, B=B, v_cap_cols=400, M=Mc)
    assert q["ret"] == 0
    A = harness.csr_to_dense(rp, ci, v, m)
    Mx = np.diag(Mc[2]) if mass else np.eye(m)
    X = q["V"] @ q["T"] @ q["V"].T
    R = A @ X @ Mx.T + Mx @ X @ A.T + B @ B.T
    exact = np.max(np.abs(np.linalg.eigvalsh(0.5 * (R + R.T))))
    assert abs(q["abs"][-1] - exact) < 1e-7 * exact
    assert exact < 1e-6 * np.linalg.norm(B.T @ B, 2)


def test_coefficient_space_with_restarts_matches_method_1(gpu):
    """Default restarts (every 20 iterations, compression at convergence) shrink
    the basis through the pivoted QR of the coefficient blocks: methods 1 and 2
    stay on the same path and return an orthonormal V."""
    rp, ci, v = problems.laplacian(40, 2)
    m = rp.size - 1
    P = {"Tolerance": 1e-6, "Expand size": 1, "Lanczos iterations": 10, "Orthogonalize candidates": 1}
    a = gpu.solve(rp, ci, v, 1, dict(P, **{"Residual method": 1}), v_cap_cols=300)
    b = gpu.solve(rp, ci, v, 1, dict(P, **{"Residual method": 2}), v_cap_cols=300)
    assert a["ret"] == 0 and b["ret"] == 0
    assert a["iterations"] == b["iterations"] and np.array_equal(a["space"], b["space"])
    err = np.abs(a["rel"] - b["rel"]) / a["rel"]
    assert np.max(err[:8]) < 1e-9 and np.max(err) < 1e-2
    assert np.linalg.norm(b["V"].T @ b["V"] - np.eye(b["k"])) < 1e-10
    assert harness.lowrank_rel_diff(a["V"], a["T"], b["V"], b["T"]) < 1e-6
