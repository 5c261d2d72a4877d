// TEST INFRASTRUCTURE -- NOT PART OF THE PRODUCT (see oracle_kernels.cpp).
//
// Procedural CPU restatement of RAILS::Solver::solve and its helpers
// (/root/reference/src/LyapunovSolver.hpp:100-482) over a CSR operator, with
// the StlWrapper semantics for every vector operation (random stream, CGS2,
// spectral norm).  No templates, no wrapper classes: flat arrays and explicit
// loops, so that it shares no code with the product's solver loop.
// The C++ reference ignores M; so does this restatement.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <map>
#include <string>
#include <vector>

extern "C" {
void oracle_spmm(int m, const int *rowptr, const int *colidx, const double *vals, int ncols,
                 const double *X, long long ldx, double *Y, long long ldy);
void oracle_dot(long long m, int a, const double *X, long long ldx, int b, const double *Y,
                long long ldy, double *out);
void oracle_gemm(long long m, int k, const double *V, long long ldv, const double *C, int ldc,
                 int c, double alpha, double beta, double *Out, long long ldo);
void oracle_orthogonalize(long long m, int n, double *V, long long ld, int first);
void oracle_random(long long m, int n, unsigned seed, long long row_begin, long long row_end,
                   double *X, long long ld);

void dgees_(const char *jobvs, const char *sort, int (*select)(double *, double *), const int *n,
            double *a, const int *lda, int *sdim, double *wr, double *wi, double *vs,
            const int *ldvs, double *work, const int *lwork, int *bwork, int *info);
void dtrsyl_(const char *trana, const char *tranb, const int *isgn, const int *m, const int *n,
             const double *a, const int *lda, const double *b, const int *ldb, double *c,
             const int *ldc, double *scale, int *info);
void dsyev_(const char *jobz, const char *uplo, const int *n, double *a, const int *lda,
            double *w, double *work, const int *lwork, int *info);
}

namespace
{

typedef std::vector<double> vec;

// C (m x n) = op(A) * op(B) for small column-major matrices, plain loops
void small_mm(bool ta, bool tb, int m, int n, int k, const double *A, int lda, const double *B,
              int ldb, double *C, int ldc)
{
    for (int j = 0; j < n; ++j)
        for (int i = 0; i < m; ++i)
        {
            double s = 0.0;
            for (int l = 0; l < k; ++l)
                s += (ta ? A[l + (size_t)i * lda] : A[i + (size_t)l * lda]) *
                     (tb ? B[j + (size_t)l * ldb] : B[l + (size_t)j * ldb]);
            C[i + (size_t)j * ldc] = s;
        }
}

int sym_eig(int n, double *a, int lda, double *w)
{
    if (n == 0) return 0;
    int info = 0, lwork = -1;
    double q = 0.0;
    dsyev_("V", "U", &n, a, &lda, w, &q, &lwork, &info);
    lwork = std::max(3 * n, (int)q);
    vec work(lwork);
    dsyev_("V", "U", &n, a, &lda, w, work.data(), &lwork, &info);
    return info;
}

// indices sorted by |value| descending (src/StlTools.hpp:12-30)
std::vector<int> by_magnitude(const double *v, int n)
{
    std::vector<std::pair<int, double> > t;
    for (int i = 0; i < n; ++i)
        t.push_back(std::make_pair(i, v[i]));
    std::sort(t.begin(), t.end(), [](std::pair<int, double> const &a, std::pair<int, double> const &b)
              { return std::abs(a.second) > std::abs(b.second); });
    std::vector<int> out;
    for (auto &p : t) out.push_back(p.first);
    return out;
}

// spectral norm of an m x n block: sqrt(lambda_max(X'X)) (src/StlWrapper.cpp:265-289)
double norm2(long long m, int n, const double *X, long long ld)
{
    vec g((size_t)n * n), w(n);
    oracle_dot(m, n, X, ld, n, X, ld, g.data());
    if (n == 1)
        return std::sqrt(std::abs(g[0]));
    sym_eig(n, g.data(), n, w.data());
    double mx = 0.0;
    for (int i = 0; i < n; ++i)
        mx = std::max(mx, std::sqrt(std::abs(w[i])));
    return mx;
}

}

extern "C" {

// A X + X A' + C = 0 on k x k matrices: SB03MD('C','X','N','T') followed by the
// sign flip of Solver::dense_solve (:348-365).  Bartels-Stewart: real Schur
// form, triangular Sylvester solve, back-transformation, symmetrised.
// Returns 0, or k+1 when the Schur factor has (nearly) opposite eigenvalues.
int oracle_dense_solve(int k, const double *A, int lda, const double *C, int ldc, double *X, int ldx)
{
    if (k < 1) return 0;
    vec S((size_t)k * k), U((size_t)k * k), wr(k), wi(k), t1((size_t)k * k), t2((size_t)k * k);
    for (int j = 0; j < k; ++j)
        for (int i = 0; i < k; ++i)
            S[i + (size_t)j * k] = A[i + (size_t)j * lda];
    int sdim = 0, info = 0, lwork = -1;
    double q = 0.0;
    dgees_("V", "N", nullptr, &k, S.data(), &k, &sdim, wr.data(), wi.data(), U.data(), &k, &q,
           &lwork, nullptr, &info);
    lwork = std::max(3 * k, (int)q);
    vec work(lwork);
    dgees_("V", "N", nullptr, &k, S.data(), &k, &sdim, wr.data(), wi.data(), U.data(), &k,
           work.data(), &lwork, nullptr, &info);
    if (info > 0) return info;
    small_mm(true, false, k, k, k, U.data(), k, C, ldc, t1.data(), k);
    small_mm(false, false, k, k, k, t1.data(), k, U.data(), k, t2.data(), k);
    int isgn = 1, sinfo = 0;
    double scale = 1.0;
    dtrsyl_("N", "T", &isgn, &k, &k, S.data(), &k, S.data(), &k, t2.data(), &k, &scale, &sinfo);
    small_mm(false, false, k, k, k, U.data(), k, t2.data(), k, t1.data(), k);
    small_mm(false, true, k, k, k, t1.data(), k, U.data(), k, t2.data(), k);
    for (int j = 0; j < k; ++j)
        for (int i = 0; i <= j; ++i)
        {
            // like Solver::dense_solve, `scale` (1 unless near overflow) is not applied
            double v = 0.5 * (t2[i + (size_t)j * k] + t2[j + (size_t)i * k]);
            X[i + (size_t)j * ldx] = -v;
            X[j + (size_t)i * ldx] = -v;
        }
    return sinfo == 1 ? k + 1 : 0;
}

// Residual Lanczos (:367-447).  Returns the number of steps taken; Q is
// m x (L+1), H (L+1) x (L+1) zero-initialised here, seed = the std::rand()
// draw the reference would make.
int oracle_resid_lanczos(long long m, int k, const double *V, long long ldv, const double *AV,
                         long long ldav, const double *T, int ldt, int n_b, const double *B,
                         long long ldb, int L, unsigned seed, double *Q, long long ldq,
                         double *H, int ldh)
{
    oracle_random(m, 1, seed, 0, m, Q, ldq);
    {
        double nrm = norm2(m, 1, Q, ldq);
        double s = 1.0 / nrm;
        for (long long r = 0; r < m; ++r) Q[r] *= s;
    }
    for (int j = 0; j <= L; ++j)
        for (int i = 0; i <= L; ++i)
            H[i + (size_t)j * ldh] = 0.0;

    vec Y(n_b), Z(std::max(k, 1)), TZ(std::max(k, 1));
    double alpha = 0.0, beta = 0.0;
    int it = 0;
    for (int i = 0; i < L; ++i)
    {
        const double *q = Q + (size_t)it * ldq;
        double *nx = Q + (size_t)(it + 1) * ldq;
        oracle_dot(m, n_b, B, ldb, 1, q, ldq, Y.data());                 // Y = B' q
        oracle_gemm(m, n_b, B, ldb, Y.data(), n_b, 1, 1.0, 0.0, nx, ldq); // q+ = B Y

        oracle_dot(m, k, V, ldv, 1, q, ldq, Z.data());
        small_mm(false, false, k, 1, k, T, ldt, Z.data(), k, TZ.data(), k);
        oracle_gemm(m, k, AV, ldav, TZ.data(), k, 1, 1.0, 1.0, nx, ldq);

        oracle_dot(m, k, AV, ldav, 1, q, ldq, Z.data());
        small_mm(false, false, k, 1, k, T, ldt, Z.data(), k, TZ.data(), k);
        oracle_gemm(m, k, V, ldv, TZ.data(), k, 1, 1.0, 1.0, nx, ldq);

        oracle_dot(m, 1, nx, ldq, 1, q, ldq, &alpha);
        H[it + (size_t)it * ldh] = alpha;

        // q+ -= alpha*q: the scaled copy is rounded before the subtraction
        for (long long r = 0; r < m; ++r) { double t = q[r] * alpha; nx[r] -= t; }
        if (it > 0)
        {
            const double *qm = Q + (size_t)(it - 1) * ldq;
            for (long long r = 0; r < m; ++r) { double t = qm[r] * beta; nx[r] -= t; }
        }

        beta = norm2(m, 1, nx, ldq);
        if (beta < 1e-14) { it++; break; }
        H[(it + 1) + (size_t)it * ldh] = beta;
        H[it + (size_t)(it + 1) * ldh] = beta;
        double s = 1.0 / beta;
        for (long long r = 0; r < m; ++r) nx[r] *= s;
        it++;
    }
    return it;
}

}
