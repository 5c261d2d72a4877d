"""The two algebraic facts the one-kernel Lanczos step rests on,
checked in numpy on random data -- no GPU, no library call:

  1. alpha_i = q' R q needs no pass over the rows: with g = [V AV B]' q and
     coef = [T g_AV ; T g_V ; g_B],  R q = [V AV B] coef  and  alpha = g . coef.
  2. the new Lanczos vector and the NEXT projection are row-local: computing both tile by
     tile (any tile height, ragged last tile) gives the same vector and, to rounding, the
     same projection as two full passes.
The recurrence is src/LyapunovSolver.hpp:384-434."""
import numpy as np
import pytest


def _problem(m, k, n, seed):
    rs = np.random.RandomState(seed)
    V = np.linalg.qr(rs.uniform(-1, 1, (m, k)))[0]
    AV = rs.uniform(-1, 1, (m, k))
    B = rs.uniform(-1, 1, (m, n))
    T = rs.uniform(-1, 1, (k, k))
    T = T + T.T
    q = rs.uniform(-1, 1, m)
    return V, AV, B, T, q / np.linalg.norm(q)


@pytest.mark.parametrize("m,k,n", [(500, 3, 1), (4099, 15, 1), (3000, 24, 4), (1000, 47, 2)])
def test_alpha_from_the_reduced_projection(m, k, n):
    V, AV, B, T, q = _problem(m, k, n, m + k)
    X = np.hstack([V, AV, B])
    g = X.T @ q
    coef = np.concatenate([T @ g[k:2 * k], T @ g[:k], g[2 * k:]])
    Rq = B @ (B.T @ q) + AV @ (T @ (V.T @ q)) + V @ (T @ (AV.T @ q))       # :395-403
    assert np.max(np.abs(X @ coef - Rq)) <= 1e-12 * np.max(np.abs(Rq))
    alpha_ref = Rq @ q                                                   # :406
    assert abs(g @ coef - alpha_ref) <= 1e-12 * max(abs(alpha_ref), np.linalg.norm(Rq))


@pytest.mark.parametrize("m,k,n,tile", [(1000, 5, 1, 128), (4099, 15, 1, 256), (3001, 30, 3, 512), (777, 2, 1, 1000)])
def test_update_and_next_projection_are_row_local(m, k, n, tile):
    V, AV, B, T, q = _problem(m, k, n, m + tile)
    rs = np.random.RandomState(1)
    q_prev = rs.uniform(-1, 1, m)
    beta_prev = 0.37
    X = np.hstack([V, AV, B])
    g = X.T @ q
    coef = np.concatenate([T @ g[k:2 * k], T @ g[:k], g[2 * k:]])
    alpha = g @ coef
    # two full passes: the new vector, then its projection
    w = X @ coef - alpha * q - beta_prev * q_prev
    g_next, nrm2 = X.T @ w, w @ w
    # one pass, tile by tile: each tile forms its rows of w and adds X_tile' w_tile
    w_t = np.zeros(m)
    g_t = np.zeros(2 * k + n)
    s_t = 0.0
    for r0 in range(0, m, tile):
        r1 = min(m, r0 + tile)
        Xt = X[r0:r1]
        wt = Xt @ coef - alpha * q[r0:r1] - beta_prev * q_prev[r0:r1]
        w_t[r0:r1] = wt
        g_t += Xt.T @ wt
        s_t += wt @ wt
    assert np.max(np.abs(w_t - w)) <= 4e-16 * np.max(np.abs(w)) * (2 * k + n)
    assert np.max(np.abs(g_t - g_next)) <= 1e-12 * max(np.max(np.abs(g_next)), 1.0)
    assert abs(s_t - nrm2) <= 1e-12 * nrm2


def test_cgs2_pass_is_the_same_kernel_with_other_scalars():
    """v - V c  =  (sum_j V_j (-c_j)) - alpha * (v * sc)  with alpha = -1, sc = 1: the
    step kernel with these scalars performs a Gram-Schmidt update, and its projection
    output is the coefficient vector of the second pass (cgs2_fused, lanczos.cu)."""
    rs = np.random.RandomState(5)
    m, k = 2000, 12
    V = np.linalg.qr(rs.uniform(-1, 1, (m, k)))[0]
    v = rs.uniform(-1, 1, m)
    c1 = V.T @ v
    alpha, sc = -1.0, 1.0
    x = V @ (-c1) - alpha * (v * sc)
    assert np.array_equal(x, V @ (-c1) + v)
    assert np.max(np.abs(x - (v - V @ c1))) <= 4e-16 * k
    c2 = V.T @ x
    y = V @ (-c2) + x
    assert np.max(np.abs(V.T @ y)) <= 1e-15 * np.linalg.norm(v) * 10      # twice is enough


def test_stable_radix_sort_by_column_gives_the_transposed_csr():
    """transpose.cu in numpy: histogram of the columns -> scan = row pointers of A';
    LSD radix sort of (column, position) by column, 8 bits a pass, each pass STABLE;
    gather.  Because CSR positions ascend with the row and every pass is stable, each row
    of A' lists its entries by ascending original row -- scipy's A.T.tocsr() with sorted
    indices, entry for entry."""
    import scipy.sparse as sp
    rs = np.random.RandomState(2)
    m, n = 900, 70001                                   # 17 bits of column index: three passes
    A = sp.random(m, n, density=0.003, format="csr", random_state=rs, data_rvs=lambda k: rs.uniform(-1, 1, k))
    A.sort_indices()
    rp, ci, v = A.indptr, A.indices, A.data
    nnz = ci.size
    rowptr_t = np.concatenate([[0], np.cumsum(np.bincount(ci, minlength=n))])
    keys, pos = ci.copy(), np.arange(nnz)
    bits = 1
    while (1 << bits) < n:
        bits += 1
    for shift in range(0, bits, 8):
        order = np.argsort((keys >> shift) & 255, kind="stable")
        keys, pos = keys[order], pos[order]
    rows_of_pos = np.searchsorted(rp, pos, side="right") - 1        # the binary search of the gather kernel
    At = A.T.tocsr()
    At.sort_indices()
    assert np.array_equal(rowptr_t, At.indptr)
    assert np.array_equal(rows_of_pos, At.indices)
    assert np.array_equal(v[pos], At.data)
