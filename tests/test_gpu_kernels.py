"""GPU parity tests, kernel by kernel, through the C ABI (include/rails_cuda.h)
against the CPU oracle on the same seeded inputs.

Bars: bit-exact for random() and the SpMM (both are order-defined in the
reference: the minstd stream of src/StlWrapper.cpp:414-423 and the CSR row sum
y_i = sum_j a_ij x_j); 1e-13 * (sum of |terms|) for the reductions (the GPU
sums in a different, fixed order) and the fma-accumulated V*C products.
"""
import ctypes

import numpy as np
import pytest

import harness
from rails_b200 import problems

pytestmark = pytest.mark.gpu

LL = ctypes.c_longlong


def _csr(api, rp, ci, v, m, ncols_global=None):
    A = ctypes.c_void_p()
    n = m if ncols_global is None else ncols_global
    api.check(api.cuda().rails_csr_create(m, m, n, 0, api.iptr(rp), api.iptr(ci), api.dptr(v),
                                          ctypes.byref(A)), "csr_create")
    return A


@pytest.mark.parametrize("m,n,seed", [(7, 3, 1), (1000, 1, 12345), (33, 4, 2147483647), (5, 2, 0)])
def test_random_bit_exact_vs_reference_golden(gpu, golden, m, n, seed):
    libc = ctypes.CDLL(None)
    libc.srand(ctypes.c_uint(seed))
    s = libc.rand()
    mv = gpu.MultiVector(m, n)
    gpu.check(gpu.cuda().rails_mv_random(mv.h, 0, n, 0, ctypes.c_uint(s)), "random")
    assert np.array_equal(mv.numpy(), golden["rng_%d_%d_%d" % (m, n, seed)])
    # a row block with an offset (row partition) reproduces the same rows
    half = gpu.MultiVector(m - m // 2, n)
    # the block is filled as rows [m//2, m) of the GLOBAL m x n fill
    gpu.check(gpu.cuda().rails_mv_random(half.h, 0, n, m // 2, ctypes.c_uint(s)), "random")
    assert np.array_equal(half.numpy(), golden["rng_%d_%d_%d" % (m, n, seed)][m // 2:])


def test_random_large_vs_oracle(gpu):
    m, n, s = 300001, 3, 1804289383
    mv = gpu.MultiVector(m, n)
    gpu.check(gpu.cuda().rails_mv_random(mv.h, 0, n, 0, ctypes.c_uint(s)), "random")
    a = np.zeros((m, n), order="F")
    harness.oracle().oracle_random(LL(m), n, ctypes.c_uint(s), LL(0), LL(m), harness.dptr(a), LL(m))
    assert np.array_equal(mv.numpy(), a)


@pytest.mark.parametrize("kind,g,dim", [("laplacian", 33, 2), ("convdiff", 17, 3), ("laplacian", 1, 2)])
@pytest.mark.parametrize("e", [1, 2, 3, 4, 7, 16])
def test_spmm_bit_exact(gpu, kind, g, dim, e):
    rp, ci, v = problems.build(kind, g, dim)
    m = rp.size - 1
    rs = np.random.RandomState(e)
    X = np.asfortranarray(rs.uniform(-1, 1, (m, e)))
    Yo = np.zeros((m, e), order="F")
    harness.oracle().oracle_spmm(m, harness.iptr(rp), harness.iptr(ci), harness.dptr(v), e,
                                 harness.dptr(X), LL(m), harness.dptr(Yo), LL(m))
    A = _csr(gpu, rp, ci, v, m)
    Xd, Yd = gpu.MultiVector.from_numpy(X), gpu.MultiVector(m, e + 1)
    gpu.check(gpu.cuda().rails_csr_spmm(A, 0, Xd.h, 0, e, Yd.h, 1), "spmm")
    assert np.array_equal(Yd.numpy(1, e), Yo)
    gpu.cuda().rails_csr_free(A)


def test_csr_create_rejects_bad_columns(gpu):
    rp = np.array([0, 2, 3], dtype=np.int32)
    v = np.ones(3)
    A = ctypes.c_void_p()
    for bad in ([0, 2, 1], [0, 1, -1]):
        ci = np.array(bad, dtype=np.int32)
        assert gpu.cuda().rails_csr_create(2, 2, 2, 0, gpu.iptr(rp), gpu.iptr(ci), gpu.dptr(v),
                                           ctypes.byref(A)) == 2
    ci = np.array([0, 1, 1], dtype=np.int32)
    gpu.check(gpu.cuda().rails_csr_create(2, 2, 2, 0, gpu.iptr(rp), gpu.iptr(ci), gpu.dptr(v),
                                          ctypes.byref(A)), "csr_create")
    gpu.cuda().rails_csr_free(A)


def test_large_upload_download_round_trip(gpu):
    # columns above 8 MB take the staged (pinned, multi-threaded) transfer path
    m = (40 << 20) // 8 + 13
    a = np.asfortranarray(np.random.RandomState(0).uniform(-1, 1, (m, 2)))
    mv = gpu.MultiVector.from_numpy(a)
    assert np.array_equal(mv.numpy(), a)
    assert np.array_equal(mv.numpy(1, 1), a[:, 1:2])
    mv.free()
    # many short columns, > 8 MB in all: several whole columns per pipeline piece;
    # the host block has a leading dimension larger than the row count
    m, k = 300007, 37
    a = np.asfortranarray(np.random.RandomState(1).uniform(-1, 1, (m, k)))
    mv = gpu.MultiVector.from_numpy(a)
    out = np.full((m + 5, k - 3), -3.0, order="F")
    gpu.check(gpu.cuda().rails_mv_download(mv.h, 2, k - 3, gpu.dptr(out), m + 5), "download")
    assert np.array_equal(out[:m], a[:, 2:k - 1])
    assert np.all(out[m:] == -3.0)
    # columns of ~20 MB: one column per piece, odd tail
    m, k = (20 << 20) // 8 + 77, 5
    a = np.asfortranarray(np.random.RandomState(2).uniform(-1, 1, (m, k)))
    mv2 = gpu.MultiVector.from_numpy(a)
    assert np.array_equal(mv2.numpy(), a)
    mv.free(); mv2.free()


def test_spmm_row_longer_than_the_staging_buffer(gpu):
    # one row with 7000 non-zeros (> the 6144-entry stage): whole-CTA fallback,
    # surrounded by ordinary rows and an empty block of rows
    rs = np.random.RandomState(11)
    m, n = 900, 7000
    rp = np.zeros(m + 1, dtype=np.int32)
    ci, v = [], []
    for i in range(m):
        if i == 300:
            cols = np.arange(n)
        elif 400 <= i < 700:
            cols = np.zeros(0, dtype=np.int64)
        else:
            cols = np.sort(rs.choice(n, size=rs.randint(0, 9), replace=False))
        ci.extend(cols); v.extend(rs.uniform(-1, 1, cols.size)); rp[i + 1] = len(ci)
    ci, v = np.array(ci, dtype=np.int32), np.array(v, dtype=np.float64)
    A = ctypes.c_void_p()
    gpu.check(gpu.cuda().rails_csr_create(m, m, n, 0, gpu.iptr(rp), gpu.iptr(ci), gpu.dptr(v),
                                          ctypes.byref(A)), "csr_create")
    for e in (1, 3):
        X = np.asfortranarray(rs.uniform(-1, 1, (n, e)))
        Xd, Yd = gpu.MultiVector.from_numpy(X), gpu.MultiVector(m, e)
        gpu.check(gpu.cuda().rails_mv_fill(Yd.h, 0, e, ctypes.c_double(7.0)), "fill")
        gpu.check(gpu.cuda().rails_csr_spmm(A, 0, Xd.h, 0, e, Yd.h, 0), "spmm")
        ref = np.zeros((m, e))
        for i in range(m):
            ref[i] = v[rp[i]:rp[i + 1]] @ X[ci[rp[i]:rp[i + 1]]]
        assert np.max(np.abs(Yd.numpy() - ref)) < 1e-12
        assert np.array_equal(Yd.numpy()[400:700], np.zeros((300, e)))
    gpu.cuda().rails_csr_free(A)


def test_spmm_transpose_and_irregular_rows(gpu):
    # random sparse rectangular matrix with empty rows and one very long row
    rs = np.random.RandomState(3)
    m, n = 700, 900
    dense = np.where(rs.uniform(size=(m, n)) < 0.01, rs.uniform(-1, 1, (m, n)), 0.0)
    dense[5, :] = rs.uniform(-1, 1, n)          # long row
    dense[10:20, :] = 0.0                       # empty rows
    rp = np.zeros(m + 1, dtype=np.int32)
    ci, v = [], []
    for i in range(m):
        nz = np.nonzero(dense[i])[0]
        ci.extend(nz); v.extend(dense[i, nz]); rp[i + 1] = len(ci)
    ci, v = np.array(ci, dtype=np.int32), np.array(v, dtype=np.float64)
    A = ctypes.c_void_p()
    gpu.check(gpu.cuda().rails_csr_create(m, m, n, 0, gpu.iptr(rp), gpu.iptr(ci), gpu.dptr(v),
                                          ctypes.byref(A)), "csr_create")
    X = np.asfortranarray(rs.uniform(-1, 1, (n, 3)))
    Xd, Yd = gpu.MultiVector.from_numpy(X), gpu.MultiVector(m, 3)
    gpu.check(gpu.cuda().rails_csr_spmm(A, 0, Xd.h, 0, 3, Yd.h, 0), "spmm")
    assert np.max(np.abs(Yd.numpy() - dense @ X)) < 1e-12
    Z = np.asfortranarray(rs.uniform(-1, 1, (m, 2)))
    Zd, Wd = gpu.MultiVector.from_numpy(Z), gpu.MultiVector(n, 2)
    gpu.check(gpu.cuda().rails_csr_spmm(A, 1, Zd.h, 0, 2, Wd.h, 0), "spmm^T")
    assert np.max(np.abs(Wd.numpy() - dense.T @ Z)) < 1e-12
    nrm = ctypes.c_double(0)
    gpu.check(gpu.cuda().rails_csr_norm2(A, ctypes.byref(nrm)), "norm2")
    assert abs(nrm.value - np.linalg.norm(dense, 2)) < 1e-6 * nrm.value
    gpu.cuda().rails_csr_free(A)


@pytest.mark.parametrize("shape", ["grid", "tall_dense_columns", "wide", "one_column", "empty"])
def test_transposed_operator_built_on_the_device_is_exact_and_ordered(gpu, shape):
    """a2: A' is built on the device (column histogram, scan, STABLE radix sort of the
    positions by column, gather).  Every row of A' must list its entries by ascending
    original row, so that A'x is the same sum, term by term in the same order, as a CSR
    loop over scipy's A.T.tocsr() (which sorts the same way): bit for bit, and the same
    on a rebuilt operator.  tall_dense_columns is B given as an Operator
    (src/LyapunovSolver.hpp:150,389: B'W): rows of A' with 120 000 entries."""
    import scipy.sparse as sp
    rs = np.random.RandomState(8)
    if shape == "grid":
        rp, ci, v = problems.build("convdiff", 40, 3)
        m = n = 40 ** 3
    elif shape == "tall_dense_columns":
        m, n = 120000, 5
        D = rs.uniform(-1, 1, (m, n))
        D[rs.uniform(size=(m, n)) < 0.3] = 0.0
        D[:, 2] = rs.uniform(0.5, 1, m)              # one completely dense column
        S = sp.csr_matrix(D)
        rp, ci, v = S.indptr.astype(np.int32), S.indices.astype(np.int32), S.data.copy()
    elif shape == "wide":
        m, n = 300, 70001                             # 17 bits of column index: three radix passes
        S = sp.random(m, n, density=0.002, format="csr", random_state=rs, data_rvs=lambda k: rs.uniform(-1, 1, k))
        rp, ci, v = S.indptr.astype(np.int32), S.indices.astype(np.int32), S.data.copy()
    elif shape == "one_column":
        m, n = 5000, 1
        rp = np.arange(m + 1, dtype=np.int32)
        ci, v = np.zeros(m, dtype=np.int32), rs.uniform(-1, 1, m)
    else:
        m, n = 100, 50
        rp, ci, v = np.zeros(m + 1, dtype=np.int32), np.zeros(0, dtype=np.int32), np.zeros(0)
    St = sp.csr_matrix((v, ci, rp), shape=(m, n)).T.tocsr()
    St.sort_indices()
    e = 2
    X = np.asfortranarray(rs.uniform(-1, 1, (m, e)))
    want = np.zeros((n, e), order="F")
    LL = ctypes.c_longlong
    if St.nnz:
        harness.oracle().oracle_spmm(n, harness.iptr(St.indptr.astype(np.int32)),
                                     harness.iptr(St.indices.astype(np.int32)),
                                     harness.dptr(np.ascontiguousarray(St.data)), e, harness.dptr(X), LL(m),
                                     harness.dptr(want), LL(n))
    outs = []
    for _ in range(2):                               # two operators: the build itself is reproducible
        A = ctypes.c_void_p()
        gpu.check(gpu.cuda().rails_csr_create(m, m, n, 0, gpu.iptr(rp), gpu.iptr(ci), gpu.dptr(v),
                                              ctypes.byref(A)), "csr_create")
        Xd, Yd = gpu.MultiVector.from_numpy(X), gpu.MultiVector(n, e)
        gpu.check(gpu.cuda().rails_mv_fill(Yd.h, 0, e, ctypes.c_double(3.0)), "fill")
        gpu.check(gpu.cuda().rails_csr_spmm(A, 1, Xd.h, 0, e, Yd.h, 0), "spmm^T")
        outs.append(Yd.numpy())
        gpu.cuda().rails_csr_free(A)
        Xd.free(); Yd.free()
    assert np.array_equal(outs[0], outs[1])
    if shape == "tall_dense_columns":
        # rows of A' longer than the staging buffer take the whole-CTA reduction of the
        # SpMM kernel, a different summation order from the CSR loop: rounding only
        assert np.max(np.abs(outs[0] - want)) <= 1e-13 * np.max(np.abs(want)) * 300
    else:
        assert np.array_equal(outs[0], want)


@pytest.mark.parametrize("m", [1, 2, 15, 511, 512, 513, 4099, 100003])
@pytest.mark.parametrize("a,b", [(1, 1), (5, 1), (1, 5), (3, 2), (37, 1), (1, 200), (200, 1),
                                 (16, 16), (9, 4), (130, 3)])
def test_dot_vs_oracle(gpu, m, a, b):
    rs = np.random.RandomState(m % 1000 + a * 7 + b)
    X = np.asfortranarray(rs.uniform(-1, 1, (m, a)))
    Y = np.asfortranarray(rs.uniform(-1, 1, (m, b)))
    Go = np.zeros((a, b), order="F")
    harness.oracle().oracle_dot(LL(m), a, harness.dptr(X), LL(m), b, harness.dptr(Y), LL(m),
                                harness.dptr(Go))
    Xd, Yd = gpu.MultiVector.from_numpy(X), gpu.MultiVector.from_numpy(Y)
    G = np.full((a + 2, b), 7.0, order="F")      # leading dimension larger than a
    gpu.check(gpu.cuda().rails_mv_dot(Xd.h, 0, a, Yd.h, 0, b, gpu.dptr(G), a + 2), "dot")
    bound = 1e-13 * (np.abs(X).T @ np.abs(Y) + 1.0)
    assert np.all(np.abs(G[:a] - Go) <= bound)
    assert np.all(G[a:] == 7.0)


@pytest.mark.parametrize("m", [5, 1000, 4099, 70001])
@pytest.mark.parametrize("a,b", [(9, 9), (16, 100), (100, 16), (64, 64), (33, 70), (130, 40), (200, 200)])
def test_block_dot_on_tensor_cores(gpu, m, a, b):
    """Both blocks wider than 8 columns: the DMMA kernel of csrc/gram.cu (the
    warm-start projection V'(AV), k x e blocks with e = 16, Gram matrices)."""
    rs = np.random.RandomState(m % 1000 + a * 7 + b)
    X = np.asfortranarray(rs.uniform(-1, 1, (m, a + 1)))
    Y = np.asfortranarray(rs.uniform(-1, 1, (m, b + 2)))
    Xd, Yd = gpu.MultiVector.from_numpy(X), gpu.MultiVector.from_numpy(Y)
    G = np.full((a + 3, b), 7.0, order="F")
    gpu.check(gpu.cuda().rails_mv_dot(Xd.h, 1, a, Yd.h, 2, b, gpu.dptr(G), a + 3), "dot")
    want = X[:, 1:].T @ Y[:, 2:]
    bound = 1e-13 * (np.abs(X[:, 1:]).T @ np.abs(Y[:, 2:]) + 1.0)
    assert np.all(np.abs(G[:a] - want) <= bound)
    assert np.all(G[a:] == 7.0)
    G2 = np.zeros((a, b), order="F")
    gpu.check(gpu.cuda().rails_mv_dot(Xd.h, 1, a, Yd.h, 2, b, gpu.dptr(G2), a), "dot")
    assert np.array_equal(G2, G[:a])                     # reproducible


@pytest.mark.parametrize("m", [3, 1025, 70001])
@pytest.mark.parametrize("k,c", [(4, 9), (7, 64), (64, 32), (70, 33), (200, 100), (13, 16)])
def test_block_gemm_on_tensor_cores(gpu, m, k, c):
    """More than 8 result columns: the DMMA kernel of csrc/gram.cu (restart
    compression V <- V X)."""
    rs = np.random.RandomState(k * 31 + c)
    V = np.asfortranarray(rs.uniform(-1, 1, (m, k + 1)))
    C = np.asfortranarray(rs.uniform(-1, 1, (k + 1, c)))
    O0 = np.asfortranarray(rs.uniform(-1, 1, (m, c + 1)))
    for alpha, beta in [(1.0, 0.0), (-1.0, 1.0), (0.5, 2.0)]:
        Vd, Od = gpu.MultiVector.from_numpy(V), gpu.MultiVector.from_numpy(O0)
        gpu.check(gpu.cuda().rails_mv_gemm(Vd.h, 1, k, gpu.dptr(C), k + 1, c, alpha, beta, Od.h, 1), "gemm")
        want = alpha * (V[:, 1:] @ C[:k]) + beta * O0[:, 1:]
        bound = 1e-13 * (np.abs(V[:, 1:]) @ np.abs(C[:k]) + 2 * np.abs(O0[:, 1:]) + 1.0)
        got = Od.numpy()
        assert np.all(np.abs(got[:, 1:] - want) <= bound)
        assert np.array_equal(got[:, 0], O0[:, 0])       # the column before the block is untouched


def test_fp64_peak_probe(gpu):
    fma, mma = ctypes.c_double(0), ctypes.c_double(0)
    gpu.check(gpu.cuda().rails_cuda_fp64_peak(0, ctypes.byref(fma)), "fp64_peak")
    gpu.check(gpu.cuda().rails_cuda_fp64_peak(1, ctypes.byref(mma)), "fp64_peak")
    print("FP64 peak: DFMA %.1f TFLOP/s, DMMA %.1f TFLOP/s" % (fma.value, mma.value))
    assert 5.0 < fma.value < 100.0 and 1.0 < mma.value < 200.0


def test_dot_is_deterministic(gpu):
    rs = np.random.RandomState(11)
    m = 1 << 20
    X = np.asfortranarray(rs.uniform(-1, 1, (m, 6)))
    Xd = gpu.MultiVector.from_numpy(X)
    G1, G2 = np.zeros((6, 6), order="F"), np.zeros((6, 6), order="F")
    gpu.check(gpu.cuda().rails_mv_dot(Xd.h, 0, 6, Xd.h, 0, 6, gpu.dptr(G1), 6), "dot")
    gpu.check(gpu.cuda().rails_mv_dot(Xd.h, 0, 6, Xd.h, 0, 6, gpu.dptr(G2), 6), "dot")
    assert np.array_equal(G1, G2)
    assert np.max(np.abs(G1 - X.T @ X)) < 1e-13 * m


def test_dot_dev_stays_on_device(gpu):
    rs = np.random.RandomState(2)
    m = 5000
    B = np.asfortranarray(rs.uniform(-1, 1, (m, 4)))
    q = np.asfortranarray(rs.uniform(-1, 1, (m, 1)))
    Bd, qd = gpu.MultiVector.from_numpy(B), gpu.MultiVector.from_numpy(q)
    Y = gpu.MultiVector(4, 1, replicated=True)
    gpu.check(gpu.cuda().rails_mv_dot_dev(Bd.h, 0, 4, qd.h, 0, 1, Y.h, 0), "dot_dev")
    assert np.max(np.abs(Y.numpy() - B.T @ q)) < 1e-12
    # B * Y with the coefficients taken from the device block
    out = gpu.MultiVector(m, 1)
    gpu.check(gpu.cuda().rails_mv_gemm_dev(Bd.h, 0, 4, Y.h, 0, 1, 1.0, 0.0, out.h, 0), "gemm_dev")
    assert np.max(np.abs(out.numpy() - B @ (B.T @ q))) < 1e-11


@pytest.mark.parametrize("m", [1, 3, 512, 1025, 70001])
@pytest.mark.parametrize("k,c", [(1, 1), (7, 1), (64, 1), (200, 1), (257, 1), (10, 2), (11, 3),
                                 (20, 4), (12, 8), (21, 11), (40, 30)])
def test_gemm_vs_oracle(gpu, m, k, c):
    rs = np.random.RandomState(k * 31 + c)
    V = np.asfortranarray(rs.uniform(-1, 1, (m, k)))
    C = np.asfortranarray(rs.uniform(-1, 1, (k + 1, c)))      # ldc > k
    O0 = np.asfortranarray(rs.uniform(-1, 1, (m, c)))
    for alpha, beta in [(1.0, 0.0), (-1.0, 1.0), (0.5, 1.0)]:
        Oo = O0.copy(order="F")
        harness.oracle().oracle_gemm(LL(m), k, harness.dptr(V), LL(m), harness.dptr(C), k + 1, c,
                                     ctypes.c_double(alpha), ctypes.c_double(beta),
                                     harness.dptr(Oo), LL(m))
        Vd, Od = gpu.MultiVector.from_numpy(V), gpu.MultiVector.from_numpy(O0)
        gpu.check(gpu.cuda().rails_mv_gemm(Vd.h, 0, k, gpu.dptr(C), k + 1, c, alpha, beta,
                                           Od.h, 0), "gemm")
        bound = 1e-13 * (np.abs(V) @ np.abs(C[:k]) + np.abs(O0) + 1.0)
        assert np.all(np.abs(Od.numpy() - Oo) <= bound)


def test_gemm_rejects_overlap(gpu):
    V = gpu.MultiVector(100, 8)
    C = np.zeros((4, 2), order="F")
    rc = gpu.cuda().rails_mv_gemm(V.h, 0, 4, gpu.dptr(C), 4, 2, 1.0, 0.0, V.h, 3)
    assert rc == 2
    rc = gpu.cuda().rails_mv_gemm(V.h, 0, 4, gpu.dptr(C), 4, 2, 1.0, 0.0, V.h, 4)
    assert rc == 0


def test_blas1_family(gpu):
    lib = gpu.cuda()
    rs = np.random.RandomState(4)
    for m in (1, 2, 9, 1000, 65537):
        X = np.asfortranarray(rs.uniform(-1, 1, (m, 5)))
        Y = np.asfortranarray(rs.uniform(-1, 1, (m, 5)))
        Xd, Yd = gpu.MultiVector.from_numpy(X), gpu.MultiVector.from_numpy(Y)
        gpu.check(lib.rails_mv_scal(Xd.h, 1, 3, 0.3), "scal")
        Xe = X.copy(); Xe[:, 1:4] *= 0.3
        assert np.array_equal(Xd.numpy(), Xe)
        gpu.check(lib.rails_mv_axpy(Yd.h, 0, -1.0, Xd.h, 2, 2), "axpy")
        Ye = Y.copy(); Ye[:, 0:2] -= Xe[:, 2:4]
        assert np.array_equal(Yd.numpy(), Ye)
        gpu.check(lib.rails_mv_axpy(Yd.h, 3, 0.7, Xd.h, 0, 1), "axpy")
        assert np.max(np.abs(Yd.numpy()[:, 3] - (Ye[:, 3] + 0.7 * Xe[:, 0]))) < 1e-15
        Ye = Yd.numpy()
        gpu.check(lib.rails_mv_add_scaled(Yd.h, 4, -1.0, 0.123, Xd.h, 4, 1), "add_scaled")
        assert np.array_equal(Yd.numpy()[:, 4], Ye[:, 4] - 0.123 * Xe[:, 4])
        gpu.check(lib.rails_mv_scaled_copy(Yd.h, 0, 2.5, Xd.h, 1, 2), "scaled_copy")
        assert np.array_equal(Yd.numpy()[:, 0:2], 2.5 * Xe[:, 1:3])
        gpu.check(lib.rails_mv_fill(Yd.h, 2, 2, -4.25), "fill")
        assert np.all(Yd.numpy()[:, 2:4] == -4.25)
        assert np.array_equal(Yd.numpy()[:, 4], Ye[:, 4] - 0.123 * Xe[:, 4])
        gpu.check(lib.rails_mv_copy_cols(Yd.h, 1, Xd.h, 0, 3), "copy_cols")
        assert np.array_equal(Yd.numpy()[:, 1:4], Xe[:, 0:3])
        out = ctypes.c_double(0)
        gpu.check(lib.rails_mv_norm_inf(Xd.h, 0, 5, ctypes.byref(out)), "norm_inf")
        assert out.value == harness.oracle().oracle_norm_inf(LL(m), 5, harness.dptr(np.asfortranarray(Xe)), LL(m))
        val = ctypes.c_double(0)
        gpu.check(lib.rails_mv_set(Xd.h, m - 1, 4, 3.5), "set")
        gpu.check(lib.rails_mv_get(Xd.h, m - 1, 4, ctypes.byref(val)), "get")
        assert val.value == 3.5


def test_argument_errors_are_reported_not_fatal(gpu):
    lib = gpu.cuda()
    X = gpu.MultiVector(10, 2)
    Y = gpu.MultiVector(11, 2)
    assert lib.rails_mv_axpy(Y.h, 0, 1.0, X.h, 0, 1) == 2          # row mismatch
    assert lib.rails_mv_fill(X.h, 1, 2, 0.0) == 2                  # range
    G = np.zeros((2, 2), order="F")
    assert lib.rails_mv_dot(X.h, 0, 2, Y.h, 0, 2, gpu.dptr(G), 2) == 2
    assert b"rows" in lib.rails_cuda_last_error()
    # empty shapes are fine
    E = gpu.MultiVector(0, 3)
    assert lib.rails_mv_fill(E.h, 0, 3, 1.0) == 0
    assert lib.rails_mv_dot(E.h, 0, 3, E.h, 0, 3, gpu.dptr(np.zeros((3, 3), order="F")), 3) == 0


def test_cgs2_known_answer_and_golden(gpu, golden):
    lib = gpu.cuda()
    # test/GenericMultiVectorWrapper_test.cpp:270-311
    X = np.zeros((10, 2), order="F")
    X[0, 0], X[0, 1], X[1, 1] = 2.3, 5.3, 2.7
    Xd = gpu.MultiVector.from_numpy(X)
    gpu.check(lib.rails_mv_orthogonalize(Xd.h, 0, 0, 1), "orthogonalize")
    E = np.zeros((10, 2)); E[0, 0] = E[1, 1] = 1.0
    assert np.max(np.abs(Xd.numpy() - E)) <= 4 * np.finfo(float).eps
    # reference CGS2 output on a random 50 x 6 block
    Xd = gpu.MultiVector.from_numpy(golden["cgs2_in"])
    gpu.check(lib.rails_mv_orthogonalize(Xd.h, 0, 0, 5), "orthogonalize")
    Q = Xd.numpy()
    assert np.max(np.abs(Q - golden["cgs2_out"])) < 1e-14
    assert np.max(np.abs(Q.T @ Q - np.eye(6))) < 1e-14


@pytest.mark.parametrize("m,k,first", [(100003, 12, 0), (100003, 40, 37), (2000, 130, 120)])
def test_cgs2_vs_oracle(gpu, m, k, first):
    rs = np.random.RandomState(k)
    X = np.asfortranarray(rs.uniform(-1, 1, (m, k)))
    if first:
        X[:, :first] = np.linalg.qr(X[:, :first])[0]
    Xo = X.copy(order="F")
    harness.oracle().oracle_orthogonalize(LL(m), k, harness.dptr(Xo), LL(m), first)
    Xd = gpu.MultiVector.from_numpy(X)
    gpu.check(gpu.cuda().rails_mv_orthogonalize(Xd.h, 0, first, k - 1), "orthogonalize")
    Q = Xd.numpy()
    assert np.max(np.abs(Q - Xo)) < 1e-12
    assert np.max(np.abs(Q.T @ Q - np.eye(k))) < 1e-13
    assert np.array_equal(Q[:, :first], X[:, :first])


@pytest.mark.parametrize("first,dups", [(20, [(22, 21)]), (9, [(10, 9), (12, 9)]), (0, [(3, 1)])])
def test_cgs2_with_duplicated_new_columns(gpu, first, dups):
    """Two of the new columns are copies of one another (what the expansion gets
    when two Ritz vectors are ghosts of the same eigenvector): the second one
    collapses to rounding noise.  Whatever direction it ends up with, the result
    must be orthonormal, and the old columns untouched."""
    m, k = 5000, first + 6
    rs = np.random.RandomState(first + 1)
    X = np.asfortranarray(rs.uniform(-1, 1, (m, k)))
    if first:
        X[:, :first] = np.linalg.qr(X[:, :first])[0]
    for dst, src in dups:
        X[:, dst] = X[:, src] * 1.5
    Xd = gpu.MultiVector.from_numpy(X)
    gpu.check(gpu.cuda().rails_mv_orthogonalize(Xd.h, 0, first, k - 1), "orthogonalize")
    Q = Xd.numpy()
    assert np.all(np.isfinite(Q))
    assert np.max(np.abs(Q.T @ Q - np.eye(k))) < 1e-9
    assert np.array_equal(Q[:, :first], X[:, :first])


def test_linearity_at_full_size(gpu):
    """Size-independent properties at the C3 row count (m = 256^3), where the
    oracle is too slow to be the checker: A(x + 2y) = Ax + 2Ay on the 3-D
    Laplacian, <x, A y> = <A x, y> (symmetry), V'(V c) = (V'V) c."""
    lib = gpu.cuda()
    g = 256
    rp, ci, v = problems.laplacian(g, 3)
    m = rp.size - 1
    A = _csr(gpu, rp, ci, v, m)
    X = gpu.MultiVector(m, 4)
    gpu.check(lib.rails_mv_random(X.h, 0, 2, 0, ctypes.c_uint(99)), "random")
    gpu.check(lib.rails_mv_copy_cols(X.h, 2, X.h, 0, 1), "copy")
    gpu.check(lib.rails_mv_axpy(X.h, 2, 2.0, X.h, 1, 1), "axpy")          # col2 = x + 2y
    Y = gpu.MultiVector(m, 4)
    gpu.check(lib.rails_csr_spmm(A, 0, X.h, 0, 3, Y.h, 0), "spmm")
    gpu.check(lib.rails_mv_axpy(Y.h, 2, -1.0, Y.h, 0, 1), "axpy")
    gpu.check(lib.rails_mv_axpy(Y.h, 2, -2.0, Y.h, 1, 1), "axpy")
    out = ctypes.c_double(0)
    gpu.check(lib.rails_mv_norm_inf(Y.h, 2, 1, ctypes.byref(out)), "norm_inf")
    assert out.value < 1e-13 * 12 * 3
    G = np.zeros((2, 2), order="F")
    gpu.check(lib.rails_mv_dot(X.h, 0, 2, Y.h, 0, 2, gpu.dptr(G), 2), "dot")
    assert abs(G[0, 1] - G[1, 0]) < 1e-12 * m
    # one row of A x checked on the host: interior row
    r = g * g * 5 + g * 7 + 9
    xs = X.numpy(0, 1)[:, 0]
    want = -6 * xs[r] + xs[r - 1] + xs[r + 1] + xs[r - g] + xs[r + g] + xs[r - g * g] + xs[r + g * g]
    got = ctypes.c_double(0)
    gpu.check(lib.rails_mv_get(Y.h, r, 0, ctypes.byref(got)), "get")
    assert abs(got.value - want) < 1e-14
    lib.rails_csr_free(A)


@pytest.mark.parametrize("case", ["lap3d_mid", "lap3d_first", "convdiff_last", "random", "no_ghosts", "empty_rank",
                                  "all_ghost"])
def test_device_halo_plan_equals_the_host_plan(gpu, case):
    """rails_csr_create builds the halo plan of a row-partitioned operator on the
    device; the host restatement rails_csr_halo_plan (what Epetra's column map +
    Import hold, src/Epetra_OperatorWrapper.cpp:87; pinned by
    tests/test_partition_gloo.py) must give the same renumbering, ghost list,
    owner rows and per-rank counts -- bit for bit."""
    import ctypes
    lib = gpu.cuda()
    rs = np.random.RandomState(5)
    world = 4
    if case.startswith("lap3d") or case.startswith("convdiff"):
        g = 24
        m = g ** 3
        rank = {"lap3d_mid": 1, "lap3d_first": 0, "convdiff_last": 3}[case]
        r0, r1 = problems.partition_rows(m, world, rank)
        rp, ci, _ = problems.build("convdiff" if case.startswith("conv") else "laplacian", g, 3, r0, r1)
    elif case == "random":
        m, world = 50021, 7
        r0, r1 = problems.partition_rows(m, world, 3)
        per = rs.randint(0, 9, r1 - r0)
        rp = np.concatenate([[0], np.cumsum(per)]).astype(np.int32)
        ci = rs.randint(0, m, int(rp[-1])).astype(np.int32)        # unsorted, duplicates, any owner
    elif case == "no_ghosts":
        m = 4096
        r0, r1 = problems.partition_rows(m, world, 2)
        rp = np.arange(r1 - r0 + 1, dtype=np.int32)
        ci = np.arange(r0, r1, dtype=np.int32)
    elif case == "empty_rank":
        m, world = 3, 4                                           # rank 3 owns no rows
        r0, r1 = problems.partition_rows(m, world, 3)
        assert r1 == r0
        rp, ci = np.zeros(1, dtype=np.int32), np.zeros(0, dtype=np.int32)
    else:                                                          # every column lives elsewhere
        m = 10000
        r0, r1 = problems.partition_rows(m, world, 1)
        rp = (2 * np.arange(r1 - r0 + 1)).astype(np.int32)
        ci = np.concatenate([rs.randint(0, r0, r1 - r0), rs.randint(r1, m, r1 - r0)]).astype(np.int32)
    rows, nnz = r1 - r0, int(rp[-1])
    ci = np.ascontiguousarray(ci)

    def plan(fn):
        lcol = np.full(max(nnz, 1), -1, dtype=np.int32)
        gg = np.full(max(nnz, 1), -1, dtype=np.int32)
        gl = np.full(max(nnz, 1), -1, dtype=np.int32)
        need = np.full(world, -1, dtype=np.int32)
        ng = ctypes.c_int(-1)
        assert fn(rows, m, r0, world, gpu.iptr(rp), gpu.iptr(ci), gpu.iptr(lcol), ctypes.byref(ng), gpu.iptr(gg),
                  gpu.iptr(gl), gpu.iptr(need)) == 0
        return lcol[:nnz], ng.value, gg[:ng.value], gl[:ng.value], need

    host, dev = plan(lib.rails_csr_halo_plan), plan(lib.rails_csr_halo_plan_device)
    assert host[1] == dev[1]
    for a, b in zip(host, dev):
        assert np.array_equal(a, b)
    if case == "no_ghosts" or case == "empty_rank":
        assert dev[1] == 0
    # an index outside the global range is an argument error on both
    if nnz:
        bad = ci.copy()
        bad[nnz // 2] = m
        for fn in (lib.rails_csr_halo_plan, lib.rails_csr_halo_plan_device):
            lcol = np.zeros(nnz, dtype=np.int32)
            need = np.zeros(world, dtype=np.int32)
            ng = ctypes.c_int(0)
            assert fn(rows, m, r0, world, gpu.iptr(rp), gpu.iptr(bad), gpu.iptr(lcol), ctypes.byref(ng), None, None,
                      gpu.iptr(need)) == 2
