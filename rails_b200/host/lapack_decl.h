// Fortran LAPACK/BLAS entry points used by the host side of rails_b200.
//
// Only k x k (projected) problems ever reach these routines; nothing with an
// m-sized dimension is sent to the host.  Provider on this image: the LP64
// OpenBLAS 0.3.15 bundled with opencv-python-headless (see build.py).
#ifndef RAILS_B200_LAPACK_DECL_H
#define RAILS_B200_LAPACK_DECL_H

extern "C" {
void dgees_(const char *jobvs, const char *sort, int (*select)(double *, double *),
            const int *n, double *a, const int *lda, int *sdim, double *wr,
            double *wi, double *vs, const int *ldvs, double *work,
            const int *lwork, int *bwork, int *info);
void dtrsyl_(const char *trana, const char *tranb, const int *isgn,
             const int *m, const int *n, const double *a, const int *lda,
             const double *b, const int *ldb, double *c, const int *ldc,
             double *scale, int *info);
void dsyev_(char *jobz, char *uplo, int *n, double *a, int *lda, double *w,
            double *work, int *ldwork, int *info);
void dsteqr_(char *compz, int *n, double *d, double *e, double *z, int *ldz,
             double *work, int *info);
void dgemm_(char *transa, char *transb, int *m, int *n, int *k, double *alpha,
            double const *a, int *lda, double const *b, int *ldb, double *beta,
            double *c, int *ldc);
}

#endif
