// TEST INFRASTRUCTURE -- NOT PART OF THE PRODUCT.
//
// CPU restatement of the arithmetic on the RAILS hot path, one plain function
// per reference method, on raw column-major arrays (leading dimension = rows).
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may load this library; the product never does.
//
// Parity status: pinned.  tests/test_oracle.py checks these functions and
// oracle_solve (oracle_solver.cpp) against the reference itself compiled here
// (oracle/_ref/librails_ref_stl.so: unmodified src/LyapunovSolver.hpp +
// src/StlWrapper.cpp) and against the reference's own known answers
// (test/SlicotWrapper_test.cpp:7-38, test/LyapunovSolverEpetra_test.cpp:50-240,
// test/GenericMultiVectorWrapper_test.cpp:270-362); the committed fixtures
// under tests/golden/ carry the reference's outputs to the GPU box.
//
// Built with -O2 -ffp-contract=off (reference default flags, no FMA
// contraction: SURVEY F12).  Row loops are OpenMP-parallel with a static
// partition and partial sums added in thread order, so a result depends on
// the thread count but not on scheduling.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>

#include <omp.h>

extern "C" {

// y_i = sum_j a_ij x_j in CSR order, separately rounded multiply and add
// (Epetra_CrsMatrix::Apply behind src/Epetra_OperatorWrapper.cpp:75-91)
void oracle_spmm(int m, const int *rowptr, const int *colidx, const double *vals,
                 int ncols, const double *X, long long ldx, double *Y, long long ldy)
{
#pragma omp parallel for schedule(static)
    for (int i = 0; i < m; ++i)
        for (int c = 0; c < ncols; ++c)
        {
            double s = 0.0;
            for (int p = rowptr[i]; p < rowptr[i + 1]; ++p)
                s += vals[p] * X[colidx[p] + (size_t)c * ldx];
            Y[i + (size_t)c * ldy] = s;
        }
}

// out (a x b, ld = a) = X' Y      (StlWrapper::dot, src/StlWrapper.cpp:394-412)
void oracle_dot(long long m, int a, const double *X, long long ldx, int b, const double *Y,
                long long ldy, double *out)
{
    const int nt = omp_get_max_threads();
    std::vector<double> part((size_t)nt * a * b, 0.0);
#pragma omp parallel
    {
        const int t = omp_get_thread_num();
        const int T = omp_get_num_threads();
        const long long lo = m * t / T, hi = m * (t + 1) / T;
        double *p = part.data() + (size_t)t * a * b;
        for (int j = 0; j < b; ++j)
            for (int i = 0; i < a; ++i)
            {
                const double *x = X + (size_t)i * ldx, *y = Y + (size_t)j * ldy;
                double s = 0.0;
                for (long long r = lo; r < hi; ++r)
                    s += x[r] * y[r];
                p[i + (size_t)j * a] = s;
            }
    }
    for (int e = 0; e < a * b; ++e)
    {
        double s = 0.0;
        for (int t = 0; t < nt; ++t)
            s += part[(size_t)t * a * b + e];
        out[e] = s;
    }
}

// Out (m x c) = alpha * V (m x k) * C (k x c, ldc) + beta * Out
// (StlWrapper::operator*, src/StlWrapper.cpp:168-187, then += / -= :147-161)
void oracle_gemm(long long m, int k, const double *V, long long ldv, const double *C, int ldc,
                 int c, double alpha, double beta, double *Out, long long ldo)
{
#pragma omp parallel for schedule(static)
    for (long long r = 0; r < m; ++r)
        for (int t = 0; t < c; ++t)
        {
            double s = 0.0;
            for (int j = 0; j < k; ++j)
                s += V[r + (size_t)j * ldv] * C[j + (size_t)t * ldc];
            double &o = Out[r + (size_t)t * ldo];
            o = (beta == 0.0) ? alpha * s : beta * o + alpha * s;
        }
}

// column-wise CGS2 of columns [first, n) (StlWrapper::orthogonalize :305-321)
void oracle_orthogonalize(long long m, int n, double *V, long long ld, int first)
{
    std::vector<double> h(std::max(n, 1));
    for (int i = first; i < n; ++i)
    {
        double *v = V + (size_t)i * ld;
        double d;
        oracle_dot(m, 1, v, ld, 1, v, ld, &d);
        double s = 1.0 / std::sqrt(std::abs(d));
        for (long long r = 0; r < m; ++r) v[r] *= s;
        if (i)
            for (int pass = 0; pass < 2; ++pass)
            {
                oracle_dot(m, i, V, ld, 1, v, ld, h.data());
                oracle_gemm(m, i, V, ld, h.data(), i, 1, -1.0, 1.0, v, ld);
            }
        oracle_dot(m, 1, v, ld, 1, v, ld, &d);
        s = 1.0 / std::sqrt(std::abs(d));
        for (long long r = 0; r < m; ++r) v[r] *= s;
    }
}

// StlWrapper::random (src/StlWrapper.cpp:414-423) with an explicit seed in
// place of std::rand(): fresh default_random_engine, row-major fill.
// Rows [row_begin, row_end) of the global m x n block are written (ld = local
// rows); the engine is advanced past the skipped rows.
void oracle_random(long long m, int n, unsigned seed, long long row_begin, long long row_end,
                   double *X, long long ld)
{
    std::default_random_engine generator(seed);
    std::uniform_real_distribution<double> distribution(-1, 1);
    (void)m;
    generator.discard(2ull * (unsigned long long)row_begin * (unsigned long long)n);
    for (long long i = row_begin; i < row_end; ++i)
        for (int j = 0; j < n; ++j)
            X[(i - row_begin) + (size_t)j * ld] = distribution(generator);
}

double oracle_norm_inf(long long m, int n, const double *X, long long ld)
{
    double out = 0.0;
    for (long long i = 0; i < m; ++i)
    {
        double s = 0.0;
        for (int j = 0; j < n; ++j)
            s += std::abs(X[i + (size_t)j * ld]);
        out = std::max(out, s);
    }
    return out;
}

int oracle_num_threads(void) { return omp_get_max_threads(); }
// The partial sums of oracle_dot are combined in thread order, so results are
// reproducible for a given thread count; tests pin it (the algorithm's path can
// depend on rounding, DESIGN.md "Reproducibility").
void oracle_set_num_threads(int n) { if (n > 0) omp_set_num_threads(n); }

}
