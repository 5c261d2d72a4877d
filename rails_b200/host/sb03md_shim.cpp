// sb03md_ stand-in with SLICOT's Fortran signature.
//
// The reference binds SLICOT's SB03MD through extern "C" sb03md_
// (/root/reference/src/SlicotWrapper.hpp:5-9) and calls it as
// sb03md('C','X','N','T', n, A, lda, X, ldx, &scale, &info)
// (/root/reference/src/SlicotWrapper.cpp:38-41 <- src/LyapunovSolver.hpp:357).
// SLICOT is a third-party Fortran library that is neither in /root/reference
// nor on this image (no version is pinned by the reference: cmake/FindSLICOT
// just finds libslicot).  This file restates the published algorithm of
// SB03MD for the one mode the solver uses (DICO='C', JOB='X', FACT='N'):
//
//     op(A)' X + X op(A) = scale * C,   op(A) = A  (TRANA='N') or A' ('T')
//
//   1. real Schur form  A = U S U'                      (dgees_)
//   2. C~ = U' C U
//   3. triangular Sylvester solve in the Schur basis    (dtrsyl_)
//   4. X = U Y U', mirrored so X is exactly symmetric like SLICOT's result
//
// On exit A holds S, U the Schur vectors, alphar/alphai the eigenvalues and
// info = n+1 when S and -S' have (nearly) common eigenvalues (dtrsyl's
// "perturbed" return), which the solver tolerates (src/LyapunovSolver.hpp:361).
// Pinned by the reference's known answers test/SlicotWrapper_test.cpp:7-38
// (see tests/test_sb03md.py) and cross-checked against scipy.
#include "lapack_decl.h"

#include <algorithm>
#include <cstring>
#include <vector>

namespace
{

// C <- op(A) * op(B) for small column-major n x n matrices.
void small_gemm(char ta, char tb, int n, const double *a, int lda,
                const double *b, int ldb, double *c, int ldc)
{
    double one = 1.0, zero = 0.0;
    dgemm_(&ta, &tb, &n, &n, &n, &one, a, &lda, b, &ldb, &zero, c, &ldc);
}

}

extern "C" void sb03md_(char *dico, char *job, char *fact, char *trans, int *n_,
                        double *A, int *lda_, double *U, int *ldu_, double *X,
                        int *ldx_, double *scale, double *sep, double *ferr,
                        double *alphar, double *alphai, int *iwork,
                        double *dwork, int *ldwork, int *info)
{
    (void)iwork; (void)dwork; (void)ldwork;
    const int n = *n_, lda = *lda_, ldu = *ldu_, ldx = *ldx_;
    *info = 0;
    *scale = 1.0;
    if (sep) *sep = 0.0;
    if (ferr) *ferr = 0.0;

    if (!(*dico == 'C' || *dico == 'c')) { *info = -1; return; }
    if (!(*job == 'X' || *job == 'x')) { *info = -2; return; }
    const bool factored = (*fact == 'F' || *fact == 'f');
    const bool transposed = (*trans == 'T' || *trans == 't' ||
                             *trans == 'C' || *trans == 'c');
    if (n < 0) { *info = -5; return; }
    if (lda < std::max(1, n)) { *info = -7; return; }
    if (ldu < std::max(1, n)) { *info = -9; return; }
    if (ldx < std::max(1, n)) { *info = -11; return; }
    if (n == 0) return;

    if (!factored)
    {
        int sdim = 0, lwork = -1, ginfo = 0;
        double wq = 0.0;
        dgees_("V", "N", nullptr, &n, A, &lda, &sdim, alphar, alphai, U, &ldu,
               &wq, &lwork, nullptr, &ginfo);
        lwork = std::max(3 * n, (int)wq);
        std::vector<double> work(lwork);
        dgees_("V", "N", nullptr, &n, A, &lda, &sdim, alphar, alphai, U, &ldu,
               work.data(), &lwork, nullptr, &ginfo);
        if (ginfo > 0) { *info = ginfo; return; }
    }

    // C~ = U' C U (contiguous n x n temporaries)
    std::vector<double> t1((size_t)n * n), t2((size_t)n * n);
    small_gemm('T', 'N', n, U, ldu, X, ldx, t1.data(), n);
    small_gemm('N', 'N', n, t1.data(), n, U, ldu, t2.data(), n);

    // TRANA='T':  A X + X A' = scale C  ->  S Y + Y S' = scale C~
    // TRANA='N':  A' X + X A = scale C  ->  S' Y + Y S = scale C~
    int isgn = 1, sinfo = 0;
    dtrsyl_(transposed ? "N" : "T", transposed ? "T" : "N", &isgn, &n, &n, A,
            &lda, A, &lda, t2.data(), &n, scale, &sinfo);
    if (sinfo == 1) *info = n + 1;

    // X = U Y U'
    small_gemm('N', 'N', n, U, ldu, t2.data(), n, t1.data(), n);
    small_gemm('N', 'T', n, t1.data(), n, U, ldu, t2.data(), n);
    for (int j = 0; j < n; ++j)
        for (int i = 0; i <= j; ++i)
        {
            double v = 0.5 * (t2[i + (size_t)j * n] + t2[j + (size_t)i * n]);
            X[i + (size_t)j * ldx] = v;
            X[j + (size_t)i * ldx] = v;
        }
}
