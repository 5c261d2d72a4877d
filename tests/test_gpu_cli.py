"""GPU tests of the command-line driver and the solution post-processing
(SURVEY.md §8(f) N1, N2): rails_cuda_main reads A.mtx / B.mtx / M.mtx and a
parameter file, solves on the GPU, writes V.mtx / T.mtx and prints the dominant
eigenvalues of X = V T V' with their share of the trace -- the job of the
reference's src/main.cpp:41-178.

Checks: the files it writes hold the bits rails_host_solve returns for the same
input; the explicit residual meets the bar of the reference's MATLAB tests
(matlab/test/test_Laplace.m:31-59: ||A X + X A' + B B'|| < tol ||B'B||); the
eigenvalues it prints are those of the dense X."""
import os
import re
import subprocess

import numpy as np
import pytest
import scipy.sparse as sp

import harness
from rails_b200 import api, problems

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "rails_b200", "rails_cuda_main")

XML = """<ParameterList name="RAILS">
  <ParameterList name="Lyapunov Solver">
    <Parameter name="Tolerance" type="double" value="1e-6"/>
    <Parameter name="Expand size" type="int" value="2"/>
    <Parameter name="Lanczos iterations" type="int" value="12"/>
    <Parameter name="Maximum iterations" type="int" value="400"/>
  </ParameterList>
  <ParameterList name="Eigenvalue Solver">
    <Parameter name="Number of Eigenvalues" type="int" value="6"/>
    <Parameter name="Convergence Tolerance" type="double" value="1e-10"/>
  </ParameterList>
</ParameterList>
"""
PARAMS = {"Tolerance": 1e-6, "Expand size": 2, "Lanczos iterations": 12, "Maximum iterations": 400}


def _write_problem(d, g=20, n_b=2, mass=None):
    rp, ci, v = problems.laplacian(g, 2)
    m = g * g
    api.mm_write_crs(os.path.join(d, "A.mtx"), rp, ci, v, m)
    B = np.random.RandomState(7).uniform(-1, 1, (m, n_b))
    Bs = sp.csr_matrix(B)
    api.mm_write_crs(os.path.join(d, "B.mtx"), Bs.indptr.astype(np.int32), Bs.indices.astype(np.int32),
                     Bs.data, n_b)
    if mass is not None:
        api.mm_write_crs(os.path.join(d, "M.mtx"), *mass, m)
    open(os.path.join(d, "params.xml"), "w").write(XML)
    return rp, ci, v, B


def _run(d, *args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([EXE, *args, "--dir", d], stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                          text=True, timeout=600, env=e)


def _printed_eigenvalues(stdout):
    lines = stdout.split("Eigenvalues of X")[1].splitlines()[1:]
    vals = []
    for line in lines:
        parts = line.split()
        if len(parts) != 2:
            break
        vals.append((float(parts[0]), float(parts[1])))
    return np.array(vals)


def test_cli_standard_equation(gpu, tmp_path):
    d = str(tmp_path)
    rp, ci, v, B = _write_problem(d)
    r = _run(d, os.path.join(d, "params.xml"), "--write-eigs")
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    for word in ("Loading matrices", "Creating solver", "Performing solve", "Profile"):
        assert word in r.stdout
    V = api.mm_read_dense(os.path.join(d, "V.mtx"))
    T = api.mm_read_dense(os.path.join(d, "T.mtx"))
    assert V.shape[0] == 400 and T.shape == (V.shape[1], V.shape[1])

    # same input through the library entry point: same bits
    o = gpu.solve(rp, ci, v, 2, PARAMS, B=B, v_cap_cols=400)
    assert o["ret"] == 0 and o["k"] == V.shape[1]
    assert np.array_equal(o["V"], V) and np.array_equal(o["T"], T)

    # explicit residual (test_Laplace.m:31-59)
    A = harness.csr_to_dense(rp, ci, v, 400)
    X = V @ T @ V.T
    R = A @ X + X @ A.T + B @ B.T
    assert np.linalg.norm(R, 2) < 1e-6 * np.linalg.norm(B.T @ B, 2) * 10

    # the printed eigenvalues are those of X, largest first, with trace shares
    lam = np.linalg.eigvalsh(X)
    lam = lam[np.argsort(-np.abs(lam))][:6]
    shown = _printed_eigenvalues(r.stdout)
    assert shown.shape == (6, 2)
    assert np.allclose(shown[:, 0], lam, rtol=1e-5, atol=0)        # printed with 6 digits
    assert np.allclose(shown[:, 1], lam / np.trace(X), rtol=1e-5, atol=0)
    # eigenvectors written with --write-eigs
    Q = api.mm_read_dense(os.path.join(d, "eigenvectors.mtx"))
    ev = api.mm_read_dense(os.path.join(d, "eigenvalues.mtx"))[:, 0]
    assert np.allclose(ev, lam, rtol=1e-10)
    assert np.linalg.norm(Q.T @ Q - np.eye(6)) < 1e-10
    assert np.linalg.norm(X @ Q - Q * ev) < 1e-10 * abs(ev[0])

    # --only-eigenvalues reads V.mtx / T.mtx back (src/main.cpp:155-164)
    r2 = _run(d, os.path.join(d, "params.xml"), "--only-eigenvalues")
    assert r2.returncode == 0, r2.stderr[-2000:]
    assert "Performing solve" not in r2.stdout
    assert np.allclose(_printed_eigenvalues(r2.stdout), shown, rtol=1e-10)


def test_cli_b_as_operator_and_ini(gpu, tmp_path):
    d = str(tmp_path)
    rp, ci, v, B = _write_problem(d)
    open(os.path.join(d, "p.ini"), "w").write(
        "[Lyapunov Solver]\nTolerance = 1e-6\nExpand size = 2\nLanczos iterations = 12\n"
        "Maximum iterations = 400\nVerbose = false\n")
    r = _run(d, os.path.join(d, "p.ini"), "--b-operator")
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "Iteration 1." not in r.stdout                     # Verbose = false
    V = api.mm_read_dense(os.path.join(d, "V.mtx"))
    T = api.mm_read_dense(os.path.join(d, "T.mtx"))
    A = harness.csr_to_dense(rp, ci, v, 400)
    X = V @ T @ V.T
    R = A @ X + X @ A.T + B @ B.T
    assert np.linalg.norm(R, 2) < 1e-5 * np.linalg.norm(B.T @ B, 2)


def test_cli_mass_matrix(gpu, tmp_path):
    d = str(tmp_path)
    mass = problems.lumped_mass(400)
    rp, ci, v, B = _write_problem(d, mass=mass)
    r = _run(d, os.path.join(d, "params.xml"))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    V = api.mm_read_dense(os.path.join(d, "V.mtx"))
    T = api.mm_read_dense(os.path.join(d, "T.mtx"))
    A = harness.csr_to_dense(rp, ci, v, 400)
    M = np.diag(mass[2])
    X = V @ T @ V.T
    R = A @ X @ M.T + M @ X @ A.T + B @ B.T
    assert np.linalg.norm(R, 2) < 1e-5 * np.linalg.norm(B.T @ B, 2)
    # an identity M.mtx selects the standard branch: same bits as without M
    api.mm_write_crs(os.path.join(d, "M.mtx"), np.arange(401, dtype=np.int32),
                     np.arange(400, dtype=np.int32), np.ones(400), 400)
    r = _run(d, os.path.join(d, "params.xml"))
    assert r.returncode == 0
    o = gpu.solve(rp, ci, v, 2, PARAMS, B=B, v_cap_cols=400)
    assert np.array_equal(o["V"], api.mm_read_dense(os.path.join(d, "V.mtx")))


def test_cli_singular_mass_and_bad_input(gpu, tmp_path):
    d = str(tmp_path)
    rp, ci, v, B = _write_problem(d)
    dm = np.ones(400)
    dm[100:140] = 0.0
    keep = dm != 0
    rpm = np.concatenate([[0], np.cumsum(keep)]).astype(np.int32)
    api.mm_write_crs(os.path.join(d, "M.mtx"), rpm, np.arange(400, dtype=np.int32)[keep], dm[keep], 400)
    # a singular M is reduced to its Schur complement (tests/test_gpu_schur.py) --
    # on one GPU and with a dense B only
    r = _run(d, os.path.join(d, "params.xml"), "--b-operator")
    assert r.returncode == 4 and "Schur" in r.stderr
    r = _run(d, os.path.join(d, "params.xml"))
    assert r.returncode == 0 and "Singular part: 40 rows, Schur complement: 360 rows" in r.stdout
    assert api.mm_read_dense(os.path.join(d, "V.mtx")).shape[0] == 400
    os.remove(os.path.join(d, "M.mtx"))
    os.remove(os.path.join(d, "B.mtx"))
    r = _run(d, os.path.join(d, "params.xml"))
    assert r.returncode == 2 and "cannot open" in r.stderr
    # Lanczos iterations <= Expand size is rejected by set_parameters (:89-95)
    _write_problem(d)
    open(os.path.join(d, "bad.ini"), "w").write("[Lyapunov Solver]\nExpand size = 5\nLanczos iterations = 5\n")
    r = _run(d, os.path.join(d, "bad.ini"))
    assert r.returncode == 2 and "Lanczos iterations" in r.stderr


def test_cli_warm_start_from_files(gpu, tmp_path):
    """Continuation (test/LyapunovSolver_test.cpp:312-352): the second run reads
    V.mtx of the first with "Restart from solution" and converges at once."""
    d = str(tmp_path)
    _write_problem(d)
    r = _run(d, os.path.join(d, "params.xml"))
    assert r.returncode == 0
    first = int(re.search(r"Iterations: (\d+)", r.stdout).group(1))
    open(os.path.join(d, "warm.xml"), "w").write(
        XML.replace('<Parameter name="Tolerance"',
                    '<Parameter name="Restart from solution" type="bool" value="true"/>\n'
                    '    <Parameter name="Tolerance"'))
    r = _run(d, os.path.join(d, "warm.xml"))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    again = int(re.search(r"Iterations: (\d+)", r.stdout).group(1))
    assert again <= 5 < first


def test_solution_eigs_against_numpy(gpu):
    rs = np.random.RandomState(11)
    m, k = 20000, 14
    V, _ = np.linalg.qr(rs.standard_normal((m, k)))
    T = rs.standard_normal((k, k))
    T = T + T.T
    lam, Q, tr = gpu.solution_eigs(V, T, num=5, tol=1e-12)
    w, Y = np.linalg.eigh(T)
    order = np.argsort(-np.abs(w))[:5]
    assert np.allclose(lam, w[order], rtol=1e-12)
    ref = V @ Y[:, order]
    sign = np.sign(np.sum(ref * Q, axis=0))
    assert np.max(np.abs(Q - ref * sign)) < 1e-12
    assert abs(tr - np.trace(T)) < 1e-10 * abs(np.trace(T)) + 1e-12
    # tolerance cut: eigenvalues below tol end the list (Epetra_OperatorWrapper.cpp:200-212)
    Tsmall = Y @ np.diag(np.where(np.abs(w) >= np.sort(np.abs(w))[-3], w, 1e-9)) @ Y.T
    lam2, Q2, _ = gpu.solution_eigs(V, Tsmall, num=10, tol=1e-6)
    assert lam2.size == 3 and Q2.shape == (m, 3)


def test_solution_eigs_non_orthonormal_basis(gpu):
    """An M-orthonormal or unnormalised V: eigenpairs of X through the Gram
    matrix, checked against the dense X."""
    rs = np.random.RandomState(12)
    m, k = 300, 7
    V = rs.standard_normal((m, k))
    T = rs.standard_normal((k, k))
    T = T + T.T
    X = V @ T @ V.T
    lam, Q, tr = gpu.solution_eigs(V, T, num=k, tol=1e-12)
    w = np.linalg.eigvalsh(X)
    w = w[np.argsort(-np.abs(w))][:k]
    assert np.allclose(lam, w, rtol=1e-10)
    assert np.linalg.norm(Q.T @ Q - np.eye(k)) < 1e-10
    assert np.linalg.norm(X @ Q - Q * lam) < 1e-9 * abs(lam[0])
    assert abs(tr - np.trace(X)) < 1e-10 * abs(np.trace(X))
