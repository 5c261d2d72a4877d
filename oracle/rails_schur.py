"""TEST INFRASTRUCTURE -- numpy/scipy restatement of the reference's
Schur-complement reduction for a singular mass matrix,
/root/reference/matlab/RAILSschur.m (the MATLAB twin of
/root/reference/src/SchurOperator.cpp:51-219).

Used by tests/ only, as the checker of rails_b200/host/CudaSchurOperator.cpp and
the device LU solves behind it.  The product never imports this file.

PARITY UNPINNED beyond inequalities: the reference factorises with KLU / MATLAB's
lu (UMFPACK), neither of which is on this image; here scipy's SuperLU stands in.
The reference's own tests of this path hold inequalities only
(matlab/test/test_MOC.m:16-34: both explicit residuals < 1e-3 on the ocean-model
fixture), and tests/test_schur_oracle.py pins this restatement to exactly those.
"""
import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spl


class Schur:
    """[S, MS, BS, Sinv, Vtrans] = RAILSschur(A, M, B, factorize), :1-92."""

    def __init__(self, A, M, B, zero_tolerance=1e-12):
        A = sp.csr_matrix(A)
        n = A.shape[0]
        d = np.asarray(M.diagonal() if sp.issparse(M) else np.diag(M) if np.ndim(M) == 2 else M)
        self.n = n
        self.idx1 = np.where(np.abs(d) < zero_tolerance)[0]          # :23
        self.idx2 = np.where(np.abs(d) > zero_tolerance)[0]          # :24
        i1, i2 = self.idx1, self.idx2
        self.A = A
        self.A11 = sp.csc_matrix(A[i1][:, i1])                       # :26-29
        self.A12 = sp.csr_matrix(A[i1][:, i2])
        self.A21 = sp.csr_matrix(A[i2][:, i1])
        self.A22 = sp.csr_matrix(A[i2][:, i2])
        self.lu11 = spl.splu(self.A11) if i1.size else None          # :31
        self.lu = None
        self.MS = np.diag(d[i2])                                     # :38-41
        B = np.asarray(B.todense() if sp.issparse(B) else B, dtype=float).reshape(n, -1)
        self.BS = B[i2, :]                                           # :43
        self.B_in_singular_part = False
        if i1.size and np.max(np.abs(B[i1, :]), initial=0.0) > np.sqrt(np.finfo(float).eps):   # :45-48
            self.BS = self.restrict(B)
            self.B_in_singular_part = True

    def A11inv(self, x):
        return self.lu11.solve(np.asarray(x, dtype=float)) if self.lu11 is not None else x

    def A11invt(self, x):
        return self.lu11.solve(np.asarray(x, dtype=float), trans="T") if self.lu11 is not None else x

    def S(self, x, transp=False):
        """AS / ASt, :35-37."""
        x = np.asarray(x, dtype=float)
        if self.idx1.size == 0:
            return (self.A22.T if transp else self.A22) @ x
        if transp:
            return self.A22.T @ x - self.A12.T @ self.A11invt(self.A21.T @ x)
        return self.A22 @ x - self.A21 @ self.A11inv(self.A12 @ x)

    def Sinv(self, x):
        """X2(A \\ Xreorder([0; x])), :60-66."""
        if self.lu is None:
            self.lu = spl.splu(sp.csc_matrix(self.A))
        full = np.zeros((self.n,) + np.shape(x)[1:])
        full[self.idx2] = x
        return self.lu.solve(full)[self.idx2]

    def restrict(self, x):
        """:70-72"""
        return x[self.idx2] - self.A21 @ self.A11inv(x[self.idx1])

    def prolongate(self, x):
        """:74-76"""
        out = np.zeros((self.n,) + np.shape(x)[1:])
        out[self.idx2] = x
        if self.idx1.size:
            out[self.idx1] = -self.A11inv(self.A12 @ x)
        return out

    def Vtrans(self, x):
        """:78-86"""
        if np.shape(x)[0] == self.n:
            return self.restrict(x)
        return self.prolongate(x)


def add_border(A, mdiag, B, border):
    """matlab/test/test_MOC.m:137-160 for arbitrary border columns."""
    A = sp.csr_matrix(A)
    n, nb = A.shape[0], border.shape[1]
    Bd = sp.csr_matrix(border)
    A2 = sp.bmat([[A, Bd], [Bd.T, sp.csr_matrix((nb, nb))]], format="csr")
    m2 = np.concatenate([mdiag, np.zeros(nb)])
    B2 = np.vstack([np.asarray(B, dtype=float).reshape(n, -1), np.zeros((nb, np.reshape(B, (n, -1)).shape[1]))])
    return A2, m2, B2
