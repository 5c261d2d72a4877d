"""Synthetic inputs for the RAILS hot path.

The reference has no C++ generators; MATLAB's `laplacian2`
(/root/reference/matlab/test/test_Laplace.m:14-21) fixes the sign/scale
convention used here: diagonal -2*dim, neighbours +1, no 1/h^2 scaling.
All matrices are CSR with int32 indices (the reference uses `int`
everywhere) and columns sorted inside each row.

Row numbering: x fastest, r = i + g*(j + g*l).  A contiguous block of rows is
therefore a slab along the slowest index, which is what the Epetra-style
linear row map hands each GPU.
"""
import numpy as np


def _stencil_csr(g, dim, diag, lo, hi, row_begin=0, row_end=None):
    """CSR rows [row_begin, row_end) of a constant-coefficient stencil.

    lo[d] / hi[d] are the coefficients of the neighbour at -1 / +1 along
    dimension d (Dirichlet boundaries: missing neighbours are dropped).
    Column indices are GLOBAL.  Returns (rowptr, colidx, vals).
    """
    m = g ** dim
    if row_end is None:
        row_end = m
    rows = np.arange(row_begin, row_end, dtype=np.int64)
    nloc = rows.size
    coords = []
    r = rows.copy()
    for _ in range(dim):
        coords.append(r % g)
        r //= g
    # candidate entries in ascending column order:
    # -g^(dim-1) ... -g, -1, 0, +1, +g ... +g^(dim-1)
    offs, coefs, valid = [], [], []
    for d in reversed(range(dim)):
        offs.append(-(g ** d)); coefs.append(lo[d]); valid.append(coords[d] > 0)
    offs.append(0); coefs.append(diag); valid.append(np.ones(nloc, dtype=bool))
    for d in range(dim):
        offs.append(g ** d); coefs.append(hi[d]); valid.append(coords[d] < g - 1)
    valid = np.stack(valid, axis=1)                       # nloc x nst
    counts = valid.sum(axis=1)
    rowptr = np.zeros(nloc + 1, dtype=np.int64)
    np.cumsum(counts, out=rowptr[1:])
    cols = rows[:, None] + np.asarray(offs, dtype=np.int64)[None, :]
    vals = np.broadcast_to(np.asarray(coefs, dtype=np.float64)[None, :], cols.shape)
    colidx = cols[valid].astype(np.int32)
    v = np.ascontiguousarray(vals[valid], dtype=np.float64)
    if rowptr[-1] >= 2 ** 31:
        raise ValueError("nnz exceeds int32 range for one partition")
    return rowptr.astype(np.int32), colidx, v


def laplacian(g, dim, row_begin=0, row_end=None):
    """2-D 5-point / 3-D 7-point Laplacian, diag -2*dim, neighbours +1."""
    return _stencil_csr(g, dim, -2.0 * dim, [1.0] * dim, [1.0] * dim,
                        row_begin, row_end)


def convection_diffusion(g, dim=3, peclet=0.5, row_begin=0, row_end=None):
    """Laplacian + first-order upwind convection, velocity (1,..,1).

    diag = -2*dim - dim*peclet, upwind (-1) neighbours +1+peclet, downwind +1
    (SURVEY.md §8(d)): nonsymmetric, spectrum in the open left half-plane.
    """
    return _stencil_csr(g, dim, -2.0 * dim - dim * peclet,
                        [1.0 + peclet] * dim, [1.0] * dim, row_begin, row_end)


def shift_diagonal(rowptr, colidx, vals, delta, row_begin=0):
    """A + delta*I on a row block (continuation step of config C5)."""
    out = vals.copy()
    nloc = rowptr.size - 1
    rows = np.repeat(np.arange(row_begin, row_begin + nloc, dtype=np.int64),
                     np.diff(rowptr))
    out[colidx.astype(np.int64) == rows] += delta
    return out


def lumped_mass(m, row_begin=0, row_end=None, seed=2):
    """Rows [row_begin, row_end) of a lumped (diagonal) mass matrix as CSR:
    d_i uniform in [0.5, 1.5], the same numbers whatever the
    partition.  Column indices are GLOBAL."""
    if row_end is None:
        row_end = m
    d = np.random.RandomState(seed).uniform(0.5, 1.5, m)[row_begin:row_end]
    n = row_end - row_begin
    return (np.arange(n + 1, dtype=np.int32), np.arange(row_begin, row_end, dtype=np.int32),
            np.ascontiguousarray(d))


_MINSTD_A, _MINSTD_P = 16807, 2147483647


def _minstd_stream(seed, start, count):
    """Draws start .. start+count-1 (0-based, the first draw is a*x0 mod p) of
    minstd_rand0 seeded with `seed`, by jump-ahead: x_j = a^(j+1) x0 mod p."""
    x0 = seed % _MINSTD_P or 1                  # linear_congruential_engine::seed
    blk = 1 << 12
    # a^1 .. a^blk mod p; products of two 31-bit numbers fit 64 bits
    pw = np.empty(blk, dtype=np.uint64)
    acc = 1
    for i in range(blk):
        acc = acc * _MINSTD_A % _MINSTD_P
        pw[i] = acc
    out = np.empty(count, dtype=np.uint64)
    P = np.uint64(_MINSTD_P)
    pos = 0
    while pos < count:
        n = min(blk, count - pos)
        base = np.uint64(x0 * pow(_MINSTD_A, start + pos, _MINSTD_P) % _MINSTD_P)
        out[pos:pos + n] = pw[:n] * base % P
        pos += n
    return out


def reference_random(m, n, seed, row_begin=0, row_end=None):
    """Rows [row_begin, row_end) of the m x n block StlWrapper::random() fills
    (/root/reference/src/StlWrapper.cpp:414-423) when std::rand() returned
    `seed`: a fresh std::default_random_engine (libstdc++: minstd_rand0), row-major
    fill from uniform_real_distribution(-1, 1), two draws per double
    (generate_canonical<double, 53>).  Same bits as the device kernel
    (rails_mv_random) and as the reference itself (tests/test_oracle.py).
    Returned column-major."""
    if row_end is None:
        row_end = m
    rows = row_end - row_begin
    x = _minstd_stream(seed, 2 * row_begin * n, 2 * rows * n).astype(np.float64)
    R = 2147483646.0                            # max - min + 1
    s = (x[0::2] - 1.0) + (x[1::2] - 1.0) * R
    u = s / (R * R)
    u[u >= 1.0] = np.nextafter(1.0, 0.0)
    return np.asfortranarray((u * 2.0 + -1.0).reshape(rows, n))


def first_rand_after_srand(seed=1):
    """std::srand(seed); std::rand() of the C library (the value that seeds the
    first random() of a reference run, test/LyapunovSolver_test.cpp:66-71)."""
    import ctypes
    libc = ctypes.CDLL(None)
    libc.srand(ctypes.c_uint(seed))
    return int(libc.rand())


def reference_B(m, n_b, row_begin=0, row_end=None):
    """`std::srand(1); B.random()` -- the B of SURVEY.md §8(d), this rank's rows.
    A solve that is handed this B passes "Skip random draws" = 1 so that V and the
    Lanczos start vectors get the std::rand() values of a reference run."""
    return reference_random(m, n_b, first_rand_after_srand(1), row_begin, row_end)


def partition_rows(m, nranks, rank):
    """Epetra_Map(m, 0, comm) linear map: first m % P ranks get one more row."""
    base, rem = divmod(m, nranks)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


CONFIGS = {
    # name: (kind, g, dim, n_b, parameter dict)   -- SURVEY.md §8(d)
    "C1": ("laplacian", 64, 2, 1, {"Tolerance": 1e-4, "Expand size": 1,
                                   "Lanczos iterations": 10}),
    "C2": ("laplacian", 1024, 2, 4, {"Tolerance": 1e-4, "Restart size": 200,
                                     "Reduced size": 100, "Expand size": 4,
                                     "Lanczos iterations": 20,
                                     "Restart iterations": -1}),
    # lumped mass matrix: the generalised branch (matlab/RAILSsolver.m:368-392)
    "C3": ("laplacian", 256, 3, 1, {"Tolerance": 1e-4, "Expand size": 1,
                                    "Lanczos iterations": 10, "Use mass matrix": 1}),
    # the same problem as the C++ reference solves it (M never read)
    "C3I": ("laplacian", 256, 3, 1, {"Tolerance": 1e-4, "Expand size": 1,
                                     "Lanczos iterations": 10}),
    # 16 new columns per iteration: the space is capped so that V and AV (134 MB
    # per column) plus the restart temporaries stay inside one GPU's 180 GB
    # BASELINE.json config 4's own list: everything else is the reference's default
    # (restart every 20 iterations, no size cap: the space reaches ~330 columns)
    "C4": ("convdiff", 256, 3, 16, {"Tolerance": 1e-4, "Expand size": 16,
                                    "Lanczos iterations": 40}),
    # round 1's variant of it with the space capped (kept for comparison with r01 numbers)
    "C4CAP": ("convdiff", 256, 3, 16, {"Tolerance": 1e-4, "Expand size": 16,
                                       "Lanczos iterations": 40, "Restart size": 192,
                                       "Reduced size": 96}),
    "C5": ("laplacian", 512, 3, 1, {"Tolerance": 1e-4, "Expand size": 1,
                                    "Lanczos iterations": 10}),
}


def build(kind, g, dim, row_begin=0, row_end=None):
    if kind == "laplacian":
        return laplacian(g, dim, row_begin, row_end)
    if kind == "convdiff":
        return convection_diffusion(g, dim, 0.5, row_begin, row_end)
    raise ValueError(kind)
