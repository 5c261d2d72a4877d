"""Generates tests/golden/moc_fixture.npz from the reference's own ocean-model
fixture, /root/reference/matlab/DataErik (Ap1.*, Bp1.co, Frcp1.co: an 8 x 8 x 4
grid with 6 unknowns per cell, n = 1536), assembled exactly as get_MOC_data and
add_border do in /root/reference/matlab/test/test_MOC.m:94-160:

    A      the Jacobian, CSR from the (beg, jco, co) triple (1-based -> 0-based)
    mdiag  the diagonal of M with everything but temperature and salinity set
           to zero (unknowns 0..3 of every cell)                     (:121-124)
    B      0.1 * F with everything but salinity set to zero          (:126-132)
    border the two checkerboard pressure modes that span the null space of A,
           as columns (the rows/columns add_border appends)          (:137-155)

Run in the build container (where /root/reference exists):
    python tests/golden/make_moc_fixture.py
The .npz travels to the GPU box, where the reference tree does not exist.
"""
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
DATA = "/root/reference/matlab/DataErik"


def main():
    beg = np.loadtxt(os.path.join(DATA, "Ap1.beg")).astype(np.int64)
    co = np.loadtxt(os.path.join(DATA, "Ap1.co"))
    info = np.loadtxt(os.path.join(DATA, "Ap1.info")).astype(np.int64)
    jco = np.loadtxt(os.path.join(DATA, "Ap1.jco")).astype(np.int64)
    mco = np.loadtxt(os.path.join(DATA, "Bp1.co"))
    frc = np.loadtxt(os.path.join(DATA, "Frcp1.co"))
    n, nnz = int(info[0]), int(info[1])
    assert n == 8 * 8 * 4 * 6 and beg.size == n + 1 and beg[-1] - 1 == nnz

    # MATLAB's sparse(i, j, v) sums duplicates and sorts; do the same
    import scipy.sparse as sp
    rows = np.repeat(np.arange(n), np.diff(beg))
    A = sp.csr_matrix((co[:nnz], (rows, jco[:nnz] - 1)), shape=(n, n))
    A.sum_duplicates()
    A.sort_indices()

    mdiag = mco.copy()
    for u in range(4):
        mdiag[u::6] = 0.0
    B = 0.1 * frc.copy()
    for u in range(5):
        B[u::6] = 0.0

    border = np.zeros((n, 2))
    for j in range(n):
        if j % 6 == 3:
            cell = j // 6
            if ((cell % 4) + ((cell // 4) % 16)) % 2 == 0:
                border[j, 0] = 1.0
            else:
                border[j, 1] = 1.0

    np.savez_compressed(os.path.join(HERE, "moc_fixture.npz"),
                        rowptr=A.indptr.astype(np.int32), colidx=A.indices.astype(np.int32),
                        vals=A.data.astype(np.float64), mdiag=mdiag, B=B.reshape(n, 1), border=border)
    print("n =", n, "nnz =", A.nnz, "singular rows =", int(np.sum(mdiag == 0.0)),
          "border columns nnz =", int(border.sum()))


if __name__ == "__main__":
    main()
