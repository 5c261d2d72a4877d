"""ctypes bindings of the product libraries (no torch, no oracle).

    cuda  -> librails_cuda.so : the C ABI of include/rails_cuda.h
    host  -> librails_host.so : rails_host_solve (solver loop over the CUDA
                                wrapper classes)

Loading fails loudly when the libraries have not been built
(`python -m rails_b200.build`); there is no Python or CPU stand-in.
"""
import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

c_double_p = ctypes.POINTER(ctypes.c_double)
c_int_p = ctypes.POINTER(ctypes.c_int)


def dptr(a):
    return None if a is None else a.ctypes.data_as(c_double_p)


def iptr(a):
    return None if a is None else a.ctypes.data_as(c_int_p)


_cuda = None
_host = None


def cuda():
    global _cuda
    if _cuda is None:
        path = os.path.join(HERE, "librails_cuda.so")
        if not os.path.exists(path):
            raise RuntimeError("librails_cuda.so is not built: run `python -m rails_b200.build`")
        lib = ctypes.CDLL(path, mode=ctypes.RTLD_GLOBAL)
        _declare_cuda(lib)
        _cuda = lib
    return _cuda


def host():
    global _host
    if _host is None:
        cuda()
        path = os.path.join(HERE, "librails_host.so")
        if not os.path.exists(path):
            raise RuntimeError("librails_host.so is not built: run `python -m rails_b200.build`")
        lib = ctypes.CDLL(path, mode=ctypes.RTLD_GLOBAL)
        lib.rails_host_last_sections.restype = ctypes.c_char_p
        lib.rails_host_last_timed_launches.restype = ctypes.c_longlong
        lib.rails_host_last_timed_syncs.restype = ctypes.c_longlong
        lib.rails_host_lu_factor.restype = ctypes.c_void_p
        lib.rails_host_schur_create.restype = ctypes.c_void_p
        lib.rails_host_session_create.restype = ctypes.c_void_p
        lib.rails_host_session_setup_seconds.restype = ctypes.c_double
        _host = lib
    return _host


# every symbol include/rails_cuda.h declares: (name, restype, argtypes)
I, LL, D, P = ctypes.c_int, ctypes.c_longlong, ctypes.c_double, ctypes.c_void_p
CUDA_SYMBOLS = [
    ("rails_cuda_init", I, [I]),
    ("rails_cuda_finalize", I, []),
    ("rails_cuda_is_initialized", I, []),
    ("rails_cuda_sync", I, []),
    ("rails_cuda_last_error", ctypes.c_char_p, []),
    ("rails_cuda_error_count", LL, []),
    ("rails_cuda_kernel_launches", LL, []),
    ("rails_cuda_host_syncs", LL, []),
    ("rails_cuda_bytes_in_use", LL, []),
    ("rails_cuda_fp64_peak", I, [I, c_double_p]),
    ("rails_cuda_event_record", I, [I]),
    ("rails_cuda_event_elapsed_ms", I, [I, I, ctypes.POINTER(ctypes.c_float)]),
    ("rails_cuda_comm_unique_id", I, [ctypes.c_char_p]),
    ("rails_cuda_comm_init", I, [I, I, ctypes.c_char_p]),
    ("rails_cuda_comm_p2p", I, []),
    ("rails_cuda_comm_rank", I, []),
    ("rails_cuda_comm_size", I, []),
    ("rails_partition_rows", I, [LL, I, I, ctypes.POINTER(LL), ctypes.POINTER(LL)]),
    ("rails_mv_alloc", I, [LL, I, I, ctypes.POINTER(P)]),
    ("rails_mv_free", I, [P]),
    ("rails_mv_rows", LL, [P]),
    ("rails_mv_ld", LL, [P]),
    ("rails_mv_cols", I, [P]),
    ("rails_mv_flags", I, [P]),
    ("rails_mv_copy_cols", I, [P, I, P, I, I]),
    ("rails_mv_upload", I, [P, I, I, c_double_p, LL]),
    ("rails_mv_download", I, [P, I, I, c_double_p, LL]),
    ("rails_mv_get", I, [P, LL, I, c_double_p]),
    ("rails_mv_set", I, [P, LL, I, D]),
    ("rails_mv_fill", I, [P, I, I, D]),
    ("rails_mv_scal", I, [P, I, I, D]),
    ("rails_mv_axpy", I, [P, I, D, P, I, I]),
    ("rails_mv_scaled_copy", I, [P, I, D, P, I, I]),
    ("rails_mv_add_scaled", I, [P, I, D, D, P, I, I]),
    ("rails_mv_norm_inf", I, [P, I, I, c_double_p]),
    ("rails_mv_dot", I, [P, I, I, P, I, I, c_double_p, I]),
    ("rails_mv_dot_batch", I, [I, ctypes.POINTER(P), c_int_p, c_int_p, ctypes.POINTER(P), c_int_p, c_int_p,
                               ctypes.POINTER(c_double_p), c_int_p]),
    ("rails_mv_dot_dev", I, [P, I, I, P, I, I, P, I]),
    ("rails_mv_dot_dev_noalloc", I, [P, I, I, P, I, I, P, I]),
    ("rails_mv_gemm", I, [P, I, I, c_double_p, I, I, D, D, P, I]),
    ("rails_mv_gemm_dev", I, [P, I, I, P, I, I, D, D, P, I]),
    ("rails_mv_multi_dot", I, [I, ctypes.POINTER(P), c_int_p, c_int_p, P, I, D, c_double_p]),
    ("rails_mv_multi_gemv", I, [I, ctypes.POINTER(P), c_int_p, c_int_p, c_double_p, P, I, P, I,
                                c_double_p]),
    ("rails_mv_lanczos_update", I, [P, I, D, P, I, D, P, I, c_double_p]),
    ("rails_mv_lanczos_run", I, [I, ctypes.POINTER(P), c_int_p, c_int_p, c_double_p, I, P, I, I,
                                 c_double_p, c_double_p, c_int_p]),
    ("rails_mv_lanczos_begin", I, [I, ctypes.POINTER(P), c_int_p, c_int_p, P, I]),
    ("rails_mv_lanczos_project", I, [I, ctypes.POINTER(P), c_int_p, c_int_p, P, I, I, I, D, c_double_p]),
    ("rails_mv_lanczos_max_columns", I, []),
    ("rails_lanczos_step_shape", I, [I, c_int_p, c_int_p]),
    ("rails_lanczos_step_timing", I, [I]),
    ("rails_lanczos_step_timing_read", I, [c_double_p, c_int_p, c_int_p]),
    ("rails_mv_orthogonalize", I, [P, I, I, I]),
    ("rails_mv_random", I, [P, I, I, LL, ctypes.c_uint]),
    ("rails_csr_create", I, [LL, LL, LL, LL, c_int_p, c_int_p, c_double_p, ctypes.POINTER(P)]),
    ("rails_csr_free", I, [P]),
    ("rails_csr_halo_plan", I, [LL, LL, LL, I, c_int_p, c_int_p, c_int_p, c_int_p, c_int_p, c_int_p,
                                c_int_p]),
    ("rails_csr_halo_plan_device", I, [LL, LL, LL, I, c_int_p, c_int_p, c_int_p, c_int_p, c_int_p, c_int_p,
                                c_int_p]),
    ("rails_csr_update_values", I, [P, c_double_p]),
    ("rails_csr_get_value", I, [P, LL, LL, c_double_p]),
    ("rails_csr_set_value", I, [P, LL, LL, D]),
    ("rails_csr_to_dense", I, [P, P, I]),
    ("rails_csr_rows", LL, [P]),
    ("rails_csr_nnz", LL, [P]),
    ("rails_csr_spmm", I, [P, I, P, I, I, P, I]),
    ("rails_csr_norm2", I, [P, c_double_p]),
    ("rails_lu_create", I, [I, c_int_p, c_int_p, c_double_p, c_int_p, c_int_p, c_double_p, c_int_p, c_int_p,
                            ctypes.POINTER(P)]),
    ("rails_lu_free", I, [P]),
    ("rails_lu_solve", I, [P, I, P, I, I, P, I]),
    ("rails_lu_nnz", LL, [P]),
    ("rails_lu_levels", I, [P, I]),
    ("rails_lu_launches_per_solve", I, [P, I]),
    ("rails_lu_plan_emulate", I, [I, c_int_p, c_int_p, c_double_p, c_int_p, c_int_p, c_double_p, c_int_p,
                                  c_int_p, I, c_double_p, I, LL, ctypes.POINTER(LL)]),
    ("rails_idx_create", I, [LL, c_int_p, ctypes.POINTER(P)]),
    ("rails_idx_free", I, [P]),
    ("rails_mv_gather_rows", I, [P, I, P, I, I, P, D]),
    ("rails_mv_scatter_rows", I, [P, I, P, I, I, P, D]),
]


def _declare_cuda(lib):
    for name, res, args in CUDA_SYMBOLS:
        f = getattr(lib, name)      # AttributeError if the symbol is missing
        f.restype = res
        f.argtypes = args


class RailsError(RuntimeError):
    pass


def check(rc, what=""):
    if rc != 0:
        raise RailsError("%s failed (%d): %s" % (what, rc, cuda().rails_cuda_last_error().decode()))


def init(device=0):
    """Select the GPU.  Raises when no CUDA device is usable."""
    check(cuda().rails_cuda_init(device), "rails_cuda_init")


class MultiVector:
    """Thin owner of a rails_mv_t for tests and the bench harness."""

    def __init__(self, rows, cols, replicated=False):
        self.h = P()
        check(cuda().rails_mv_alloc(rows, cols, 1 if replicated else 0, ctypes.byref(self.h)), "mv_alloc")
        self.rows, self.cols = rows, cols

    @classmethod
    def from_numpy(cls, a, replicated=False):
        a = np.asfortranarray(a, dtype=np.float64)
        if a.ndim == 1:
            a = a.reshape(-1, 1, order="F")
        mv = cls(a.shape[0], a.shape[1], replicated)
        if a.size:
            check(cuda().rails_mv_upload(mv.h, 0, a.shape[1], dptr(a), max(a.shape[0], 1)), "mv_upload")
        return mv

    def numpy(self, col0=0, ncols=None):
        ncols = self.cols - col0 if ncols is None else ncols
        out = np.zeros((self.rows, ncols), order="F")
        if out.size:
            check(cuda().rails_mv_download(self.h, col0, ncols, dptr(out), max(self.rows, 1)), "mv_download")
        return out

    def free(self):
        if self.h:
            cuda().rails_mv_free(self.h)
            self.h = P()

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def solve(rowptr, colidx, vals, n_b, params, B=None, seed=1, global_rows=None, row_offset=0,
          v_cap_cols=0, V0=None, time_from=0, time_to=0, hist_cap=4096, M=None, nullspace=None):
    """rails_host_solve_x through ctypes.  M = (rowptr, colidx, vals) of the mass
    matrix rows of this rank (used with params["Use mass matrix"] = 1); nullspace =
    this rank's rows of a space to project out of V.  Returns a dict."""
    lib = host()
    rows = rowptr.size - 1
    global_rows = rows if global_rows is None else global_rows
    keys = [k.encode() for k in params]
    karr = (ctypes.c_char_p * len(keys))(*keys)
    pv = np.array([float(v) for v in params.values()], dtype=np.float64)
    ha, hr = np.zeros(hist_cap), np.zeros(hist_cap)
    hk, hs = np.zeros(hist_cap, dtype=np.int32), np.zeros(hist_cap)
    hl, k = ctypes.c_int(0), ctypes.c_int(0)
    V = T = None
    if v_cap_cols > 0:
        V = np.zeros((rows, v_cap_cols), order="F")
        T = np.zeros((v_cap_cols, v_cap_cols), order="F")
        if V0 is not None:
            k.value = V0.shape[1]
            V[:, :k.value] = V0
    tm = ctypes.c_float(0)
    ss, su, sd = ctypes.c_double(0), ctypes.c_double(0), ctypes.c_double(0)
    Bc = None if B is None else np.asfortranarray(B, dtype=np.float64)
    mrp, mci, mv = M if M is not None else (None, None, None)
    Nc = None if nullspace is None else np.asfortranarray(nullspace, dtype=np.float64).reshape(rows, -1, order="F")
    ret = lib.rails_host_solve_x(
        ctypes.c_int(rows), ctypes.c_int(global_rows), ctypes.c_int(row_offset),
        iptr(rowptr), iptr(colidx), dptr(vals), iptr(mrp), iptr(mci), dptr(mv),
        ctypes.c_int(n_b), dptr(Bc), ctypes.c_uint(seed),
        ctypes.c_int(len(keys)), karr, dptr(pv), ctypes.c_int(hist_cap), dptr(ha), dptr(hr),
        iptr(hk), dptr(hs), ctypes.byref(hl), ctypes.byref(k), ctypes.c_int(v_cap_cols),
        dptr(V), dptr(T), ctypes.c_int(time_from), ctypes.c_int(time_to), ctypes.byref(tm),
        ctypes.byref(ss), ctypes.byref(su), ctypes.byref(sd),
        ctypes.c_int(0 if Nc is None else Nc.shape[1]), dptr(Nc))
    n = min(hl.value, hist_cap)
    out = dict(ret=ret, iterations=hl.value, k=k.value, abs=ha[:n].copy(), rel=hr[:n].copy(),
               space=hk[:n].copy(), seconds=hs[:n].copy(), timed_ms=tm.value,
               solve_seconds=ss.value, setup_seconds=su.value, dense_solve_seconds=sd.value,
               sections=lib.rails_host_last_sections().decode(),
               timed_launches=lib.rails_host_last_timed_launches(),
               timed_syncs=lib.rails_host_last_timed_syncs())
    if V is not None and k.value <= v_cap_cols:
        out["V"] = V[:, :k.value]
        out["T"] = T[:k.value, :k.value]
    return out


class Session:
    """rails_host_session_*: operator(s), B and the solution basis stay on the device
    between solves (continuation: update_operator + "Restart from solution" = 1)."""

    def __init__(self, rowptr, colidx, vals, n_b, B=None, seed=1, global_rows=None, row_offset=0, M=None,
                 skip_draws=0):
        lib = host()
        self.rows = rowptr.size - 1
        self.global_rows = self.rows if global_rows is None else global_rows
        Bc = None if B is None else np.asfortranarray(B, dtype=np.float64)
        mrp, mci, mv = M if M is not None else (None, None, None)
        self.h = ctypes.c_void_p(lib.rails_host_session_create(
            ctypes.c_int(self.rows), ctypes.c_int(self.global_rows), ctypes.c_int(row_offset),
            iptr(rowptr), iptr(colidx), dptr(vals), iptr(mrp), iptr(mci), dptr(mv),
            ctypes.c_int(n_b), dptr(Bc), ctypes.c_uint(seed), ctypes.c_int(skip_draws)))
        if not self.h:
            raise RailsError("rails_host_session_create failed: %s" % cuda().rails_cuda_last_error().decode())
        self.setup_seconds = lib.rails_host_session_setup_seconds(self.h)

    def update_operator(self, vals):
        v = np.ascontiguousarray(vals, dtype=np.float64)
        if host().rails_host_session_update_operator(self.h, dptr(v)):
            raise RailsError("rails_host_session_update_operator failed: %s" % cuda().rails_cuda_last_error().decode())

    def set_start(self, V0):
        V0 = np.asfortranarray(V0, dtype=np.float64)
        if host().rails_host_session_set_start(self.h, ctypes.c_int(V0.shape[1]), dptr(V0)):
            raise RailsError("rails_host_session_set_start failed")

    def solve(self, params, time_from=0, time_to=0, hist_cap=4096):
        lib = host()
        keys = [k.encode() for k in params]
        karr = (ctypes.c_char_p * len(keys))(*keys)
        pv = np.array([float(v) for v in params.values()], dtype=np.float64)
        ha, hr = np.zeros(hist_cap), np.zeros(hist_cap)
        hk, hs = np.zeros(hist_cap, dtype=np.int32), np.zeros(hist_cap)
        hl, k = ctypes.c_int(0), ctypes.c_int(0)
        tm = ctypes.c_float(0)
        ss, sd = ctypes.c_double(0), ctypes.c_double(0)
        ret = lib.rails_host_session_solve(
            self.h, ctypes.c_int(len(keys)), karr, dptr(pv), ctypes.c_int(hist_cap), dptr(ha), dptr(hr),
            iptr(hk), dptr(hs), ctypes.byref(hl), ctypes.byref(k), ctypes.c_int(time_from),
            ctypes.c_int(time_to), ctypes.byref(tm), ctypes.byref(ss), ctypes.byref(sd))
        n = min(hl.value, hist_cap)
        return dict(ret=ret, iterations=hl.value, k=k.value, abs=ha[:n].copy(), rel=hr[:n].copy(),
                    space=hk[:n].copy(), seconds=hs[:n].copy(), timed_ms=tm.value, solve_seconds=ss.value,
                    dense_solve_seconds=sd.value, sections=lib.rails_host_last_sections().decode(),
                    timed_launches=lib.rails_host_last_timed_launches(),
                    timed_syncs=lib.rails_host_last_timed_syncs())

    def download(self):
        """(V, T) of the last solution: this rank's rows of V and the k x k T."""
        k = host().rails_host_session_space_size(self.h)
        if k < 0:
            raise RailsError("the session holds no solution")
        V = np.zeros((self.rows, max(k, 1)), order="F")
        T = np.zeros((max(k, 1), max(k, 1)), order="F")
        if host().rails_host_session_download(self.h, ctypes.c_int(max(k, 1)), dptr(V), dptr(T)) < 0:
            raise RailsError("rails_host_session_download failed")
        return V[:, :k], T[:k, :k]

    def free(self):
        if self.h:
            host().rails_host_session_free(self.h)
            self.h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


# ---- MatrixMarket / parameter files / post-processing (host library) ---------
def _host_message():
    f = host().rails_host_last_message
    f.restype = ctypes.c_char_p
    return f().decode()


def mm_read_crs(path, row_begin=0, row_end=-1):
    """Rows [row_begin, row_end) of a coordinate file as (rowptr, colidx, vals,
    global_rows, cols); the reader the command-line driver uses for A, B, M."""
    lib = host()
    dims = (ctypes.c_longlong * 4)()
    if lib.rails_host_mm_read_crs(path.encode(), ctypes.c_longlong(row_begin),
                                  ctypes.c_longlong(row_end), dims):
        raise RailsError(_host_message())
    rows, global_rows, cols, nnz = (int(x) for x in dims)
    rp = np.zeros(rows + 1, dtype=np.int32)
    ci = np.zeros(nnz, dtype=np.int32)
    v = np.zeros(nnz)
    lib.rails_host_mm_fetch_crs(iptr(rp), iptr(ci), dptr(v))
    return rp, ci, v, global_rows, cols


def mm_read_dense(path):
    lib = host()
    r, c = ctypes.c_int(0), ctypes.c_int(0)
    if lib.rails_host_mm_read_dense(path.encode(), ctypes.byref(r), ctypes.byref(c)):
        raise RailsError(_host_message())
    out = np.zeros((r.value, c.value), order="F")
    if out.size:
        lib.rails_host_mm_fetch_dense(dptr(out))
    return out


def mm_write_dense(path, a, comment=""):
    a = np.asfortranarray(a, dtype=np.float64)
    if a.ndim == 1:
        a = a.reshape(-1, 1, order="F")
    if host().rails_host_mm_write_dense(path.encode(), ctypes.c_int(a.shape[0]), ctypes.c_int(a.shape[1]),
                                        dptr(a), ctypes.c_longlong(max(a.shape[0], 1)), comment.encode()):
        raise RailsError(_host_message())


def mm_write_crs(path, rowptr, colidx, vals, cols, global_rows=None, row_begin=0):
    rows = rowptr.size - 1
    global_rows = rows if global_rows is None else global_rows
    rowptr = np.ascontiguousarray(rowptr, dtype=np.int32)
    colidx = np.ascontiguousarray(colidx, dtype=np.int32)
    vals = np.ascontiguousarray(vals, dtype=np.float64)
    if host().rails_host_mm_write_crs(path.encode(), ctypes.c_int(rows), ctypes.c_int(global_rows),
                                      ctypes.c_int(cols), ctypes.c_int(row_begin), iptr(rowptr),
                                      iptr(colidx), dptr(vals)):
        raise RailsError(_host_message())


def params_get(path, sublist, key, default=None):
    """Value of `key` in `sublist` of an XML/INI parameter file as a float, or
    `default` when it is absent."""
    val, found = ctypes.c_double(0.0), ctypes.c_int(0)
    if host().rails_host_params_get(path.encode(), (sublist or "").encode(), key.encode(),
                                    ctypes.byref(val), ctypes.byref(found)):
        raise RailsError(_host_message())
    return val.value if found.value else default


def params_to_xml(path):
    buf = ctypes.create_string_buffer(1 << 16)
    if host().rails_host_params_to_xml(path.encode(), buf, ctypes.c_int(len(buf))):
        raise RailsError(_host_message())
    return buf.value.decode()


def solution_eigs(V, T, num=10, tol=1e-6, global_rows=None):
    """Dominant eigenpairs and trace of X = V T V' on the GPU
    (rails_host_solution_eigs).  Returns (eigenvalues, eigenvectors, trace)."""
    V = np.asfortranarray(V, dtype=np.float64)
    T = np.asfortranarray(T, dtype=np.float64)
    rows, k = V.shape
    global_rows = rows if global_rows is None else global_rows
    lam = np.zeros(max(k, 1))
    Q = np.zeros((rows, max(k, 1)), order="F")
    found, trace = ctypes.c_int(0), ctypes.c_double(0.0)
    rc = host().rails_host_solution_eigs(
        ctypes.c_int(rows), ctypes.c_int(global_rows), ctypes.c_int(k), dptr(V),
        ctypes.c_longlong(max(rows, 1)), dptr(T), ctypes.c_int(max(k, 1)), ctypes.c_int(num),
        ctypes.c_double(tol), ctypes.byref(found), dptr(lam), dptr(Q), ctypes.c_longlong(max(rows, 1)),
        ctypes.byref(trace))
    if rc:
        raise RailsError("rails_host_solution_eigs failed (%d): %s" % (rc, _host_message()))
    return lam[:found.value].copy(), Q[:, :found.value].copy(), trace.value


# ---- singular mass matrix: host sparse LU, Schur-complement operator -------------
def _csr32(rowptr, colidx, vals):
    return (np.ascontiguousarray(rowptr, dtype=np.int32), np.ascontiguousarray(colidx, dtype=np.int32),
            np.ascontiguousarray(vals, dtype=np.float64))


class SparseLU:
    """Host factorisation P A Q = L U (rails_b200/host/SparseLU.cpp).  solve()
    is host arithmetic (factorisation checks); device_solve() uploads the factors
    and solves on the GPU (rails_lu_create / rails_lu_solve)."""

    def __init__(self, rowptr, colidx, vals, pivot_threshold=0.1, order=None):
        rp, ci, v = _csr32(rowptr, colidx, vals)
        self.n = rp.size - 1
        info = ctypes.c_int(0)
        o = None if order is None else np.ascontiguousarray(order, dtype=np.int32)
        self.h = ctypes.c_void_p(host().rails_host_lu_factor(
            ctypes.c_int(self.n), iptr(rp), iptr(ci), dptr(v), ctypes.c_double(pivot_threshold), iptr(o),
            ctypes.byref(info)))
        self.info = info.value
        out, piv = (ctypes.c_longlong * 3)(), (ctypes.c_double * 2)()
        host().rails_host_lu_info(self.h, out, piv)
        self.nnz_l, self.nnz_u, self.off_diagonal_pivots = int(out[0]), int(out[1]), int(out[2])
        self.min_pivot, self.max_pivot = piv[0], piv[1]

    def permutations(self):
        pinv, q = np.zeros(self.n, dtype=np.int32), np.zeros(self.n, dtype=np.int32)
        host().rails_host_lu_permutations(self.h, iptr(pinv), iptr(q))
        return pinv, q

    def solve(self, b, trans=False):
        x = np.asfortranarray(np.array(b, dtype=np.float64).reshape(self.n, -1, order="F"))
        host().rails_host_lu_solve(self.h, dptr(x), ctypes.c_int(x.shape[1]), ctypes.c_longlong(self.n),
                                   ctypes.c_int(1 if trans else 0))
        return x

    def emulate_device_schedule(self, b, trans=False):
        """The solve in the order and from the arrays the kernels use, on the host
        (rails_lu_plan_emulate): checks chains, slots and windows without a GPU."""
        x = np.asfortranarray(np.array(b, dtype=np.float64).reshape(self.n, -1, order="F"))
        stats = (ctypes.c_longlong * 12)()
        rc = host().rails_host_lu_plan_emulate(self.h, ctypes.c_int(1 if trans else 0), dptr(x),
                                               ctypes.c_int(x.shape[1]), ctypes.c_longlong(self.n), stats)
        if rc:
            raise RailsError("rails_lu_plan_emulate failed (%d)" % rc)
        keys = ("levels", "segments", "chains", "windows", "chain_rows", "chain_entries")
        self.plan_stats = [dict(zip(keys, [int(stats[6 * t + i]) for i in range(6)])) for t in range(2)]
        return x

    def device_solve(self, b, trans=False, repeat=0):
        x = np.asfortranarray(np.array(b, dtype=np.float64).reshape(self.n, -1, order="F"))
        ms, stats = ctypes.c_float(0), (ctypes.c_longlong * 6)()
        rc = host().rails_host_lu_device_solve(self.h, ctypes.c_int(1 if trans else 0), dptr(x),
                                               ctypes.c_int(x.shape[1]), ctypes.c_longlong(self.n),
                                               ctypes.c_int(repeat), ctypes.byref(ms), stats)
        if rc:
            raise RailsError("rails_host_lu_device_solve failed (%d): %s"
                             % (rc, cuda().rails_cuda_last_error().decode()))
        self.device_stats = dict(nnz=int(stats[0]), levels=[int(stats[i]) for i in range(1, 5)],
                                 launches_per_solve=int(stats[5]), ms_per_solve=ms.value)
        return x

    def __del__(self):
        try:
            host().rails_host_lu_free(self.h)
        except Exception:
            pass


def minimum_degree(rowptr, colidx):
    rp = np.ascontiguousarray(rowptr, dtype=np.int32)
    ci = np.ascontiguousarray(colidx, dtype=np.int32)
    order = np.zeros(rp.size - 1, dtype=np.int32)
    host().rails_host_minimum_degree(ctypes.c_int(rp.size - 1), iptr(rp), iptr(ci), iptr(order))
    return order


def nested_dissection(rowptr, colidx, leaf=200):
    rp = np.ascontiguousarray(rowptr, dtype=np.int32)
    ci = np.ascontiguousarray(colidx, dtype=np.int32)
    order = np.zeros(rp.size - 1, dtype=np.int32)
    host().rails_host_nested_dissection(ctypes.c_int(rp.size - 1), iptr(rp), iptr(ci), ctypes.c_int(leaf),
                                        iptr(order))
    return order


class SchurOperator:
    """The Schur-complement operator on the GPU (rails_b200/host/CudaSchurOperator.cpp)
    for parity tests of its pieces: S, S', S^-1, restrict, prolongate."""

    S, ST, SINV, RESTRICT, PROLONGATE = 0, 1, 2, 3, 4

    def __init__(self, rowptr, colidx, vals, mdiag, border=None, zero_tolerance=1e-12, scale=1.0):
        rp, ci, v = _csr32(rowptr, colidx, vals)
        md = np.ascontiguousarray(mdiag, dtype=np.float64)
        bd = None if border is None else np.asfortranarray(border, dtype=np.float64)
        info = ctypes.c_int(0)
        self.h = ctypes.c_void_p(host().rails_host_schur_create(
            ctypes.c_int(rp.size - 1), iptr(rp), iptr(ci), dptr(v), dptr(md),
            ctypes.c_int(0 if bd is None else bd.shape[1]), dptr(bd), ctypes.c_double(zero_tolerance),
            ctypes.c_double(scale), ctypes.byref(info)))
        if info.value or not self.h:
            raise RailsError("rails_host_schur_create failed (%d): %s"
                             % (info.value, cuda().rails_cuda_last_error().decode()))
        nt, n1, n2 = ctypes.c_int(0), ctypes.c_int(0), ctypes.c_int(0)
        host().rails_host_schur_sizes(self.h, ctypes.byref(nt), ctypes.byref(n1), ctypes.byref(n2))
        self.n_total, self.n1, self.n2 = nt.value, n1.value, n2.value
        self.idx1, self.idx2 = np.zeros(self.n1, dtype=np.int32), np.zeros(self.n2, dtype=np.int32)
        host().rails_host_schur_indices(self.h, iptr(self.idx1), iptr(self.idx2))

    def apply(self, what, X, repeat=0):
        rows_in = self.n_total if what == self.RESTRICT else self.n2
        rows_out = self.n_total if what == self.PROLONGATE else self.n2
        X = np.asfortranarray(np.array(X, dtype=np.float64).reshape(rows_in, -1, order="F"))
        Y = np.zeros((rows_out, X.shape[1]), order="F")
        ms = ctypes.c_float(0)
        rc = host().rails_host_schur_apply(self.h, ctypes.c_int(what), dptr(X), ctypes.c_int(X.shape[1]),
                                           dptr(Y), ctypes.c_int(repeat), ctypes.byref(ms))
        if rc:
            raise RailsError("rails_host_schur_apply failed (%d): %s"
                             % (rc, cuda().rails_cuda_last_error().decode()))
        self.ms_per_apply = ms.value
        return Y

    def stats(self):
        st, fs = (ctypes.c_longlong * 6)(), ctypes.c_double(0)
        host().rails_host_schur_stats(self.h, st, ctypes.byref(fs))
        return dict(factor_nnz=int(st[0]), levels_l=int(st[1]), levels_u=int(st[2]),
                    launches_per_solve=int(st[3]), off_diagonal_pivots=int(st[4]), mvps=int(st[5]),
                    factor_seconds=fs.value)

    def free(self):
        if self.h:
            host().rails_host_schur_free(self.h)
            self.h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def solve_schur(rowptr, colidx, vals, mdiag, B, params, border=None, seed=1, v_cap_cols=400,
                hist_cap=4096, V0=None):
    """rails_host_solve_schur: A X M' + M X A' + B B' = 0 with a singular diagonal
    M through the Schur complement.  Returns a dict with V (full rows, prolongated,
    border rows included), V_reduced, T, the history and the set-up facts.  V0: a
    previous V (all rows) to warm-start from, with params["Restart from solution"] = 1."""
    lib = host()
    rp, ci, v = _csr32(rowptr, colidx, vals)
    n = rp.size - 1
    md = np.ascontiguousarray(mdiag, dtype=np.float64)
    Bc = np.asfortranarray(np.array(B, dtype=np.float64).reshape(n, -1, order="F"))
    bd = None if border is None else np.asfortranarray(border, dtype=np.float64)
    nb = 0 if bd is None else bd.shape[1]
    keys = [k.encode() for k in params]
    karr = (ctypes.c_char_p * len(keys))(*keys)
    pv = np.array([float(x) for x in params.values()], dtype=np.float64)
    ha, hr = np.zeros(hist_cap), np.zeros(hist_cap)
    hk = np.zeros(hist_cap, dtype=np.int32)
    hl, k = ctypes.c_int(0), ctypes.c_int(0)
    n2_max = int(np.sum(np.abs(md) >= params.get("Schur zero tolerance", 1e-12)))
    Vf = np.zeros((n + nb, v_cap_cols), order="F")
    Vr = np.zeros((max(n2_max, 1), v_cap_cols), order="F")
    T = np.zeros((v_cap_cols, v_cap_cols), order="F")
    sizes, stats, secs = (ctypes.c_int * 2)(), (ctypes.c_longlong * 6)(), (ctypes.c_double * 2)()
    V0c = None if V0 is None else np.asfortranarray(np.array(V0, dtype=np.float64).reshape(n + nb, -1, order="F"))
    ret = lib.rails_host_solve_schur(
        ctypes.c_int(n), iptr(rp), iptr(ci), dptr(v), dptr(md), ctypes.c_int(nb), dptr(bd),
        ctypes.c_int(Bc.shape[1]), dptr(Bc), ctypes.c_uint(seed), ctypes.c_int(len(keys)), karr, dptr(pv),
        ctypes.c_int(hist_cap), dptr(ha), dptr(hr), iptr(hk), ctypes.byref(hl), ctypes.byref(k),
        ctypes.c_int(v_cap_cols), dptr(Vf), dptr(Vr), dptr(T), sizes, stats, secs,
        ctypes.c_int(0 if V0c is None else V0c.shape[1]), dptr(V0c))
    m = min(hl.value, hist_cap)
    out = dict(ret=ret, iterations=hl.value, k=k.value, abs=ha[:m].copy(), rel=hr[:m].copy(),
               space=hk[:m].copy(), n1=sizes[0], n2=sizes[1], setup_seconds=secs[0],
               solve_seconds=secs[1],
               stats=dict(factor_nnz=int(stats[0]), levels_l=int(stats[1]), levels_u=int(stats[2]),
                          launches_per_solve=int(stats[3]), off_diagonal_pivots=int(stats[4]),
                          mvps=int(stats[5])))
    if 0 < k.value <= v_cap_cols:
        out["V"] = Vf[:, :k.value]
        out["V_reduced"] = Vr[:sizes[1], :k.value]
        out["T"] = T[:k.value, :k.value]
    return out
