"""Test-side ctypes bindings for the checkers: oracle/liboracle.so (CPU
restatement) and oracle/_ref/librails_ref_stl.so (the reference itself).
Nothing here is imported by the product."""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

c_double_p = ctypes.POINTER(ctypes.c_double)
c_int_p = ctypes.POINTER(ctypes.c_int)
LL = ctypes.c_longlong


def dptr(a):
    return None if a is None else a.ctypes.data_as(c_double_p)


def iptr(a):
    return None if a is None else a.ctypes.data_as(c_int_p)


_oracle = None
_ref = None


def oracle():
    global _oracle
    if _oracle is None:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import build_oracle
        lib = ctypes.CDLL(build_oracle.build())
        lib.oracle_norm_inf.restype = ctypes.c_double
        # machine-independent summation order in the tests (bench.py's CPU
        # baseline sets all host threads itself)
        lib.oracle_set_num_threads(int(os.environ.get("RAILS_ORACLE_THREADS", "4")))
        _oracle = lib
    return _oracle


def ref_path():
    # RAILS_REF_STL_LIB: an instrumented build of the same library
    # (tests/test_host_sanitizers.py runs the solver-loop tests against an ASan build)
    return os.environ.get("RAILS_REF_STL_LIB") or os.path.join(ROOT, "oracle", "_ref", "librails_ref_stl.so")


def ref():
    """The reference arm; None when oracle/_ref was never built here."""
    global _ref
    if _ref is None and os.path.exists(ref_path()):
        _ref = ctypes.CDLL(ref_path())
    return _ref


def _params(params):
    keys = [k.encode() for k in params]
    karr = (ctypes.c_char_p * len(keys))(*keys)
    pv = np.array([float(v) for v in params.values()], dtype=np.float64)
    return len(keys), karr, pv


def oracle_solve(rowptr, colidx, vals, n_b, params, B=None, seed=1, v_cap_cols=0, V0=None,
                 hist_cap=4096, M=None):
    lib = oracle()
    m = rowptr.size - 1
    nk, karr, pv = _params(params)
    ha, hr = np.zeros(hist_cap), np.zeros(hist_cap)
    hk, hs = np.zeros(hist_cap, dtype=np.int32), np.zeros(hist_cap)
    hl, k = ctypes.c_int(0), ctypes.c_int(0)
    V = T = None
    if v_cap_cols > 0:
        V = np.zeros((m, v_cap_cols), order="F")
        T = np.zeros((v_cap_cols, v_cap_cols), order="F")
        if V0 is not None:
            k.value = V0.shape[1]
            V[:, :k.value] = V0
    ss, sd = ctypes.c_double(0), ctypes.c_double(0)
    Bc = None if B is None else np.asfortranarray(B, dtype=np.float64)
    mrp, mci, mv = M if M is not None else (None, None, None)
    ret = lib.oracle_solve_m(
        ctypes.c_int(m), iptr(rowptr), iptr(colidx), dptr(vals), iptr(mrp), iptr(mci), dptr(mv),
        ctypes.c_int(n_b), dptr(Bc),
        ctypes.c_uint(seed), ctypes.c_int(nk), karr, dptr(pv), ctypes.c_int(hist_cap), dptr(ha),
        dptr(hr), iptr(hk), dptr(hs), ctypes.byref(hl), ctypes.byref(k), ctypes.c_int(v_cap_cols),
        dptr(V), dptr(T), ctypes.byref(ss), ctypes.byref(sd))
    n = min(hl.value, hist_cap)
    out = dict(ret=ret, iterations=hl.value, k=k.value, abs=ha[:n].copy(), rel=hr[:n].copy(),
               space=hk[:n].copy(), seconds=hs[:n].copy(), solve_seconds=ss.value,
               dense_solve_seconds=sd.value)
    if V is not None and k.value <= v_cap_cols:
        out["V"] = V[:, :k.value]
        out["T"] = T[:k.value, :k.value]
    return out


def ref_stl_solve(rowptr, colidx, vals, n_b, params, B=None, seed=1, v_cap_cols=0, hist_cap=4096,
                  entry="ref_stl_solve"):
    lib = ref()
    m = rowptr.size - 1
    nk, karr, pv = _params(params)
    ha, hr = np.zeros(hist_cap), np.zeros(hist_cap)
    hl, k = ctypes.c_int(0), ctypes.c_int(0)
    V = T = None
    if v_cap_cols > 0:
        V = np.zeros((m, v_cap_cols), order="F")
        T = np.zeros((v_cap_cols, v_cap_cols), order="F")
    ss = ctypes.c_double(0)
    Bc = None if B is None else np.asfortranarray(B, dtype=np.float64)
    ret = getattr(lib, entry)(
        ctypes.c_int(m), iptr(rowptr), iptr(colidx), dptr(vals), ctypes.c_int(n_b), dptr(Bc),
        ctypes.c_uint(seed), ctypes.c_int(nk), karr, dptr(pv), ctypes.c_int(hist_cap), dptr(ha),
        dptr(hr), ctypes.byref(hl), ctypes.byref(k), ctypes.c_int(v_cap_cols), dptr(V), dptr(T),
        ctypes.byref(ss))
    n = min(hl.value, hist_cap)
    out = dict(ret=ret, iterations=hl.value, k=k.value, abs=ha[:n].copy(), rel=hr[:n].copy(),
               solve_seconds=ss.value)
    if V is not None and k.value <= v_cap_cols:
        out["V"] = V[:, :k.value]
        out["T"] = T[:k.value, :k.value]
    return out


_ref_csr = None


def ref_csr():
    """The reference loop over its own StlWrapper + a CSR/OpenMP operator
    (oracle/_ref/librails_ref_csr.so); None when oracle/_ref was never built."""
    global _ref_csr
    path = os.path.join(ROOT, "oracle", "_ref", "librails_ref_csr.so")
    if _ref_csr is None and os.path.exists(path):
        _ref_csr = ctypes.CDLL(path)
    return _ref_csr


def ref_csr_solve(rowptr, colidx, vals, n_b, params, B=None, seed=1, skip_draws=0, v_cap_cols=0,
                  hist_cap=4096):
    """Sparse problems of any size through the UNMODIFIED reference loop on the CPU.
    `seconds` holds the time at which each iteration's progress line was printed."""
    lib = ref_csr()
    m = rowptr.size - 1
    nk, karr, pv = _params(params)
    ha, hr, hs = np.zeros(hist_cap), np.zeros(hist_cap), np.zeros(hist_cap)
    hl, k = ctypes.c_int(0), ctypes.c_int(0)
    V = T = None
    if v_cap_cols > 0:
        V = np.zeros((m, v_cap_cols), order="F")
        T = np.zeros((v_cap_cols, v_cap_cols), order="F")
    ss = ctypes.c_double(0)
    Bc = None if B is None else np.asfortranarray(B, dtype=np.float64)
    ret = lib.ref_csr_solve(
        ctypes.c_int(m), iptr(rowptr), iptr(colidx), dptr(vals), ctypes.c_int(n_b), dptr(Bc),
        ctypes.c_uint(seed), ctypes.c_int(skip_draws), ctypes.c_int(nk), karr, dptr(pv),
        ctypes.c_int(hist_cap), dptr(ha), dptr(hr), dptr(hs), ctypes.byref(hl), ctypes.byref(k),
        ctypes.c_int(v_cap_cols), dptr(V), dptr(T), ctypes.byref(ss))
    n = min(hl.value, hist_cap)
    out = dict(ret=ret, iterations=hl.value, k=k.value, abs=ha[:n].copy(), rel=hr[:n].copy(),
               seconds=hs[:n].copy(), solve_seconds=ss.value)
    if V is not None and k.value <= v_cap_cols:
        out["V"] = V[:, :k.value]
        out["T"] = T[:k.value, :k.value]
    return out


def lowrank_rel_diff(V1, T1, V2, T2):
    """||V1 T1 V1' - V2 T2 V2'||_F / ||V2 T2 V2'||_F through the factors."""
    G11, G22, G12 = V1.T @ V1, V2.T @ V2, V1.T @ V2
    n1 = np.trace(T1 @ G11 @ T1.T @ G11)
    n2 = np.trace(T2 @ G22 @ T2.T @ G22)
    cross = np.trace(T1 @ G12 @ T2.T @ G12.T)
    return np.sqrt(max(n1 - 2 * cross + n2, 0.0) / n2)


def csr_to_dense(rowptr, colidx, vals, n):
    m = rowptr.size - 1
    A = np.zeros((m, n))
    for i in range(m):
        A[i, colidx[rowptr[i]:rowptr[i + 1]]] = vals[rowptr[i]:rowptr[i + 1]]
    return A


def assert_same_solve(d, o, tol, early=8, early_tol=1e-8, late_tol=1e-4, x_tol=1e-6):
    """Parity of two runs of the same solve (dicts from *_solve with V, T).

    The reference algorithm is only reproducible to rounding while its
    un-reorthogonalised Lanczos stays inside the range of the residual operator
    (rank 2k+n >= L); past a breakdown the expansion vector is picked by
    rounding noise and a solve settles on one of a few trajectories -- the CPU
    oracle itself needs 85 or 98 iterations on the 2-D Laplacian 128^2 depending
    on its OpenMP thread count (DESIGN.md "Reproducibility").  So: when the two
    runs took the same number of iterations they must agree to rounding
    (history and X); when they did not, both must have converged to the same X
    within the tolerance they were asked for, in a comparable number of
    iterations.  Returns True when the strict branch was taken."""
    assert d["ret"] == o["ret"]
    if (d["iterations"], d["k"]) == (o["iterations"], o["k"]):
        rel = np.abs(d["rel"] - o["rel"]) / o["rel"]
        assert np.max(rel[:early]) < early_tol, rel[:early]
        assert np.max(rel) < late_tol, np.max(rel)
        assert lowrank_rel_diff(d["V"], d["T"], o["V"], o["T"]) < x_tol
        return True
    assert abs(d["iterations"] - o["iterations"]) <= max(2, 0.2 * o["iterations"]), \
        (d["iterations"], o["iterations"])
    assert np.abs(d["rel"][0] - o["rel"][0]) / o["rel"][0] < 1e-8
    # two converged runs differ by the residual tolerance times the conditioning
    # of the Lyapunov operator (measured on the CPU oracle alone: 8.5e-4 between
    # its 50- and 54-iteration runs on the 2-D Laplacian 32^2 at tol 1e-4)
    dx = lowrank_rel_diff(d["V"], d["T"], o["V"], o["T"])
    assert dx < 300 * tol, dx
    return False


def stl_solve_ext(rowptr, colidx, vals, B, params, M=None, V0=None, nullspace=None, seed=1,
                  v_cap_cols=300, hist_cap=4096, inverse=None):
    """The PRODUCT's solver loop over the reference's StlWrapper (CPU, dense),
    with a dense mass matrix, a warm-start basis, a null space and a dense A^-1
    (for "Projection method" > 1)."""
    lib = ref()
    m = rowptr.size - 1
    if inverse is not None:
        inv = np.asfortranarray(inverse, dtype=np.float64)
        lib.b200loop_stl_set_inverse(ctypes.c_int(m), dptr(inv))
    nk, karr, pv = _params(params)
    B = np.asfortranarray(B, dtype=np.float64).reshape(m, -1, order="F")
    Mc = None if M is None else np.asfortranarray(M, dtype=np.float64)
    Vc = None if V0 is None else np.asfortranarray(V0, dtype=np.float64).reshape(m, -1, order="F")
    Nc = None if nullspace is None else np.asfortranarray(nullspace, dtype=np.float64).reshape(m, -1, order="F")
    ha, hr = np.zeros(hist_cap), np.zeros(hist_cap)
    hk, hres = np.zeros(hist_cap, dtype=np.int32), np.zeros(hist_cap, dtype=np.int32)
    hl, k = ctypes.c_int(0), ctypes.c_int(0)
    V = np.zeros((m, v_cap_cols), order="F")
    T = np.zeros((v_cap_cols, v_cap_cols), order="F")
    ret = lib.b200loop_stl_solve_ext(
        ctypes.c_int(m), iptr(rowptr), iptr(colidx), dptr(vals), dptr(Mc), ctypes.c_int(B.shape[1]),
        dptr(B), ctypes.c_uint(seed), ctypes.c_int(nk), karr, dptr(pv),
        ctypes.c_int(0 if Vc is None else Vc.shape[1]), dptr(Vc),
        ctypes.c_int(0 if Nc is None else Nc.shape[1]), dptr(Nc),
        ctypes.c_int(hist_cap), dptr(ha), dptr(hr), iptr(hk), iptr(hres), ctypes.byref(hl),
        ctypes.byref(k), ctypes.c_int(v_cap_cols), dptr(V), dptr(T))
    n = min(hl.value, hist_cap)
    out = dict(ret=ret, iterations=hl.value, k=k.value, abs=ha[:n].copy(), rel=hr[:n].copy(),
               space=hk[:n].copy(), restarted=hres[:n].copy())
    if k.value <= v_cap_cols:
        out["V"] = V[:, :k.value]
        out["T"] = T[:k.value, :k.value]
    return out
