// TEST INFRASTRUCTURE (oracle/_ref build): the UNMODIFIED reference solver
// loop (/root/reference/src/LyapunovSolver.hpp) over the reference's own
// StlWrapper backend (/root/reference/src/StlWrapper.cpp), behind a C entry
// point so pytest / bench.py can drive it with ctypes.
//
// Compiled by oracle/build_ref.py from the sources where they lie under
// /root/reference; only the resulting .so (oracle/_ref/) travels.  The Stl
// backend is dense-only (A is an m x m column-major array and A*W is dgemm,
// src/StlWrapper.cpp:168-187), so this arm is usable up to m ~ 1e4.
#include <algorithm>
#include <chrono>
#include <iostream>
#include <utility>

#include "capture.hpp"

#include "src/LyapunovSolver.hpp"
#include "src/StlWrapper.hpp"

using namespace RAILS;

extern "C" int ref_stl_solve(
    int m, const int *rowptr, const int *colidx, const double *vals,
    int n_b, const double *B_in, unsigned seed,
    int nparams, const char *const *keys, const double *pvals,
    int hist_cap, double *hist_abs, double *hist_rel, int *hist_len,
    int *k_out, int v_cap_cols, double *V_out, double *T_out,
    double *seconds)
{
    std::srand(seed);

    StlWrapper A(m, m);
    A = 0.0;
    for (int i = 0; i < m; ++i)
        for (int p = rowptr[i]; p < rowptr[i + 1]; ++p)
            A(i, colidx[p]) = vals[p];

    StlWrapper B(m, n_b);
    if (B_in)
    {
        for (int j = 0; j < n_b; ++j)
            for (int i = 0; i < m; ++i)
                B(i, j) = B_in[i + (size_t)j * m];
    }
    else
        B.random();

    // M is stored and never read by the C++ solver; the
    // reference's own tests pass B for it (test/LyapunovSolver_test.cpp:138).
    Solver<StlWrapper, StlWrapper, StlWrapper> solver(A, B, B);
    rails_oracle::ParameterList params(nparams, keys, pvals);
    int ret = solver.set_parameters(params);
    if (ret)
        return 100 + ret;

    StlWrapper V(m, 1);
    StlWrapper T;

    std::string log;
    {
        rails_oracle::CoutCapture cap;
        auto t0 = std::chrono::steady_clock::now();
        ret = solver.solve(V, T);
        auto t1 = std::chrono::steady_clock::now();
        if (seconds)
            *seconds = std::chrono::duration<double>(t1 - t0).count();
        log = cap.str();
    }

    std::vector<double> ha, hr;
    int len = rails_oracle::parse_history(log, ha, hr);
    if (hist_len)
        *hist_len = len;
    for (int i = 0; i < std::min(len, hist_cap); ++i)
    {
        if (hist_abs) hist_abs[i] = ha[i];
        if (hist_rel) hist_rel[i] = hr[i];
    }

    int k = V.N();
    if (k_out)
        *k_out = k;
    if (V_out && T_out && k <= v_cap_cols)
    {
        for (int j = 0; j < k; ++j)
            for (int i = 0; i < m; ++i)
                V_out[i + (size_t)j * m] = V(i, j);
        for (int j = 0; j < k; ++j)
            for (int i = 0; i < k; ++i)
                T_out[i + (size_t)j * v_cap_cols] = T(i, j);
    }
    return ret;
}

// Direct access to the reference's StlWrapper::random() stream
// (src/StlWrapper.cpp:414-423) for pinning the device RNG.
extern "C" void ref_stl_random(int m, int n, unsigned seed, double *out)
{
    std::srand(seed);
    StlWrapper X(m, n);
    X.random();
    for (int j = 0; j < n; ++j)
        for (int i = 0; i < m; ++i)
            out[i + (size_t)j * m] = X(i, j);
}

// Reference SlicotWrapper entry (src/SlicotWrapper.cpp:8-49) over the
// sb03md_ stand-in, for the known answers of test/SlicotWrapper_test.cpp.
extern "C" int ref_sb03md(int n, double *A, int lda, double *X, int ldx,
                          double *scale)
{
    int info = 0;
    sb03md('C', 'X', 'N', 'T', n, A, lda, X, ldx, scale, &info);
    return info;
}

// Reference CGS2 (src/StlWrapper.cpp:305-321) on a caller-supplied block.
extern "C" void ref_stl_orthogonalize(int m, int n, double *X)
{
    StlWrapper V(m, n);
    for (int j = 0; j < n; ++j)
        for (int i = 0; i < m; ++i)
            V(i, j) = X[i + (size_t)j * m];
    V.orthogonalize();
    for (int j = 0; j < n; ++j)
        for (int i = 0; i < m; ++i)
            X[i + (size_t)j * m] = V(i, j);
}

// ---------------------------------------------------------------------------
// The PRODUCT's solver loop (rails_b200/host/RailsSolver.hpp) instantiated
// with the reference's StlWrapper for all three concepts.  Same backend, same
// operation order => must reproduce ref_stl_solve bit for bit; this is how the
// shipped loop is pinned to the reference loop (tests/test_solver_loop.py).
#include "RailsSolver.hpp"

extern "C" int b200loop_stl_solve(
    int m, const int *rowptr, const int *colidx, const double *vals,
    int n_b, const double *B_in, unsigned seed,
    int nparams, const char *const *keys, const double *pvals,
    int hist_cap, double *hist_abs, double *hist_rel, int *hist_len,
    int *k_out, int v_cap_cols, double *V_out, double *T_out,
    double *seconds)
{
    std::srand(seed);

    StlWrapper A(m, m);
    A = 0.0;
    for (int i = 0; i < m; ++i)
        for (int p = rowptr[i]; p < rowptr[i + 1]; ++p)
            A(i, colidx[p]) = vals[p];

    StlWrapper B(m, n_b);
    if (B_in)
    {
        for (int j = 0; j < n_b; ++j)
            for (int i = 0; i < m; ++i)
                B(i, j) = B_in[i + (size_t)j * m];
    }
    else
        B.random();

    RAILS::b200::Solver<StlWrapper, StlWrapper, StlWrapper> solver(A, B, B);
    rails_oracle::ParameterList params(nparams, keys, pvals);
    int ret = solver.set_parameters(params);
    if (ret)
        return 100 + ret;
    solver.set_verbose(false);

    StlWrapper V(m, 1);
    StlWrapper T;
    auto t0 = std::chrono::steady_clock::now();
    ret = solver.solve(V, T);
    if (seconds)
        *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();

    auto const &h = solver.history();
    if (hist_len)
        *hist_len = (int)h.size();
    for (int i = 0; i < std::min((int)h.size(), hist_cap); ++i)
    {
        if (hist_abs) hist_abs[i] = h[i].residual;
        if (hist_rel) hist_rel[i] = h[i].relative;
    }
    int k = V.N();
    if (k_out)
        *k_out = k;
    if (V_out && T_out && k <= v_cap_cols)
    {
        for (int j = 0; j < k; ++j)
            for (int i = 0; i < m; ++i)
                V_out[i + (size_t)j * m] = V(i, j);
        for (int j = 0; j < k; ++j)
            for (int i = 0; i < k; ++i)
                T_out[i + (size_t)j * v_cap_cols] = T(i, j);
    }
    return ret;
}

// Dense A^-1 handed to the NEXT b200loop_stl_solve_ext call (opts.Ainv of
// matlab/RAILSsolver.m:19-24; "Projection method" > 1)
static std::vector<double> g_pending_inverse;
extern "C" void b200loop_stl_set_inverse(int m, const double *Ainv)
{
    g_pending_inverse.assign(Ainv, Ainv + (size_t)m * m);
}

// The product loop over StlWrapper with everything the MATLAB-feature options
// need (tests/test_matlab_features.py): a dense mass matrix, a warm-start
// basis V0 and a null space.  M_in / V0 / N_in may be NULL.
extern "C" int b200loop_stl_solve_ext(
    int m, const int *rowptr, const int *colidx, const double *vals,
    const double *M_in, int n_b, const double *B_in, unsigned seed,
    int nparams, const char *const *keys, const double *pvals,
    int k0, const double *V0, int n_null, const double *N_in,
    int hist_cap, double *hist_abs, double *hist_rel, int *hist_k, int *hist_restarted,
    int *hist_len, int *k_out, int v_cap_cols, double *V_out, double *T_out)
{
    std::srand(seed);

    StlWrapper A(m, m);
    A = 0.0;
    for (int i = 0; i < m; ++i)
        for (int p = rowptr[i]; p < rowptr[i + 1]; ++p)
            A(i, colidx[p]) = vals[p];

    StlWrapper M(m, m);
    M = 0.0;
    for (int j = 0; j < m; ++j)
        for (int i = 0; i < m; ++i)
            M(i, j) = M_in ? M_in[i + (size_t)j * m] : (i == j ? 1.0 : 0.0);

    StlWrapper B(m, n_b);
    for (int j = 0; j < n_b; ++j)
        for (int i = 0; i < m; ++i)
            B(i, j) = B_in[i + (size_t)j * m];

    RAILS::b200::Solver<StlWrapper, StlWrapper, StlWrapper> solver(A, B, M);
    rails_oracle::ParameterList params(nparams, keys, pvals);
    int ret = solver.set_parameters(params);
    if (ret)
        return 100 + ret;
    solver.set_verbose(false);
    if (g_pending_inverse.size() == (size_t)m * m)
    {
        StlWrapper Ainv(m, m);
        for (int j = 0; j < m; ++j)
            for (int i = 0; i < m; ++i)
                Ainv(i, j) = g_pending_inverse[i + (size_t)j * m];
        solver.set_inverse(Ainv);
    }
    g_pending_inverse.clear();
    if (n_null > 0 && N_in)
    {
        StlWrapper N(m, n_null);
        for (int j = 0; j < n_null; ++j)
            for (int i = 0; i < m; ++i)
                N(i, j) = N_in[i + (size_t)j * m];
        solver.set_nullspace(N);
    }

    StlWrapper V(m, std::max(k0, 1));
    for (int j = 0; j < k0; ++j)
        for (int i = 0; i < m; ++i)
            V(i, j) = V0[i + (size_t)j * m];
    StlWrapper T;
    ret = solver.solve(V, T);

    auto const &h = solver.history();
    if (hist_len)
        *hist_len = (int)h.size();
    for (int i = 0; i < std::min((int)h.size(), hist_cap); ++i)
    {
        if (hist_abs) hist_abs[i] = h[i].residual;
        if (hist_rel) hist_rel[i] = h[i].relative;
        if (hist_k) hist_k[i] = h[i].space_size;
        if (hist_restarted) hist_restarted[i] = h[i].restarted ? 1 : 0;
    }
    int k = V.N();
    if (k_out)
        *k_out = k;
    if (V_out && T_out && k <= v_cap_cols && T.M() >= k && T.N() >= k)
    {
        for (int j = 0; j < k; ++j)
            for (int i = 0; i < m; ++i)
                V_out[i + (size_t)j * m] = V(i, j);
        for (int j = 0; j < k; ++j)
            for (int i = 0; i < k; ++i)
                T_out[i + (size_t)j * v_cap_cols] = T(i, j);
    }
    return ret;
}
