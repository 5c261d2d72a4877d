// C entry points of librails_host.so: the solver loop over the CUDA wrappers,
// callable from Python (bench.py, tests) or any FFI with plain pointers.
//
// rails_host_solve is the user-facing call of this framework: host CSR arrays
// and (optionally) a host B go in, the low-rank factors V (m x k) and T (k x k)
// of X = V T V' come back to host memory, together with the iteration history.
#include "CudaMultiVectorWrapper.hpp"
#include "CudaOperatorWrapper.hpp"
#include "HostDenseMatrix.hpp"
#include "RailsSolver.hpp"
#include "CudaFusedLanczos.hpp"

#include <chrono>
#include <dlfcn.h>
#include <cstdlib>
#include <cstring>
#include <map>
#include <sstream>
#include <string>

using namespace RAILS;

namespace
{

// duck-typed ParameterList: template<class T> T get(name, default)
// (contract of src/LyapunovSolver.hpp:40-70)
class KeyValueParameters
{
    std::map<std::string, double> values_;

public:
    KeyValueParameters(int n, const char *const *keys, const double *vals)
    {
        for (int i = 0; i < n; ++i)
            values_[keys[i]] = vals[i];
    }
    template<typename T>
    T get(std::string const &name, T def)
    {
        auto it = values_.find(name);
        return it == values_.end() ? def : (T)it->second;
    }
};

typedef b200::Solver<CudaOperatorWrapper, CudaMultiVectorWrapper, HostDenseMatrix> CudaSolver;

std::string g_sections;

// The host BLAS only ever sees k x k matrices (k <= a few hundred): keep its
// thread pool small.  On a 200-core host OpenBLAS otherwise starts one thread
// per core the first time a call crosses its threading threshold (measured:
// a 700 ms stall in the middle of a solve).  RAILS_BLAS_THREADS overrides.
void limit_blas_threads()
{
    static bool done = false;
    if (done) return;
    done = true;
    typedef void (*setter)(int);
    setter f = (setter)dlsym(RTLD_DEFAULT, "openblas_set_num_threads");
    if (!f) return;
    const char *e = std::getenv("RAILS_BLAS_THREADS");
    // one thread: the projected solve (real Schur form of a k x k matrix) is
    // level-1/2 bound; measured 195 us (1 thread) vs 330 us (4 threads) at k = 25
    int n = e ? std::atoi(e) : 1;
    if (n > 0) f(n);
}
long long g_timed_launches = 0;

}

extern "C" {

// Returns the solver's code (0 converged, -1 stopped unconverged, 1 fell out of
// the loop), 100 + code for a rejected parameter list, 200 for a CUDA failure.
//
//   rows/global_rows/row_offset : this rank's row block of the square operator
//       (single rank: rows = global_rows, row_offset = 0); colidx is GLOBAL
//   B_in   : this rank's rows of B (rows x n_b, column-major, ld = rows), or
//            NULL for B = random() after std::srand(seed)  (the reference
//            tests' way of making B, test/LyapunovSolver_test.cpp:66-71)
//   V_io   : warm start (params "Restart from solution" = 1): on entry k_io
//            columns of V (this rank's rows, ld = rows); on exit the solution
//            basis if it fits v_cap_cols columns.  T_out is v_cap_cols^2.
//   timed region: CUDA events are recorded on the library stream when
//            iteration `time_from` and iteration `time_to` finish; their
//            distance is returned in *timed_ms (0 if not reached).
//   m_rowptr/m_colidx/m_vals : this rank's rows of the mass matrix M (same row
//            block as A), or NULL.  Used only with the parameter
//            "Use mass matrix" = 1 (generalised equation A X M' + M X A' + B B' = 0,
//            matlab/RAILSsolver.m:368-392); otherwise M is stored and never read,
//            like the reference C++ loop.
//   n_null/N_in : this rank's rows of a null space to project out of every
//            vector that enters V (rows x n_null, ld = rows; opts.nullspace of
//            matlab/RAILSsolver.m:28-29), or 0/NULL
int rails_host_solve_x(int rows, int global_rows, int row_offset,
                     const int *rowptr, const int *colidx, const double *vals,
                     const int *m_rowptr, const int *m_colidx, const double *m_vals,
                     int n_b, const double *B_in, unsigned seed,
                     int nparams, const char *const *keys, const double *pvals,
                     int hist_cap, double *hist_abs, double *hist_rel, int *hist_k,
                     double *hist_sec, int *hist_len,
                     int *k_io, int v_cap_cols, double *V_io, double *T_out,
                     int time_from, int time_to, float *timed_ms,
                     double *solve_seconds, double *setup_seconds, double *dense_solve_seconds,
                     int n_null, const double *N_in)
{
    if (!rails_cuda_is_initialized())
    {
        std::cerr << "rails_host_solve: rails_cuda_init has not been called" << std::endl;
        return 200;
    }
    limit_blas_threads();
    auto t_setup = std::chrono::steady_clock::now();
    std::srand(seed);

    CudaOperatorWrapper A(rows, global_rows, global_rows, row_offset, rowptr, colidx, vals);
    if (!A.device_handle())
        return 200;

    KeyValueParameters params(nparams, keys, pvals);
    // a caller that made B itself with `srand(seed); B.random()` and passes it in
    // asks for that std::rand() value to be skipped ("Skip random draws" = 1), so
    // that V and the Lanczos start vectors get the seeds of a reference run
    for (int i = 0; i < params.get("Skip random draws", 0); ++i)
        (void)std::rand();

    CudaMultiVectorWrapper B(global_rows, n_b);
    if (B_in)
    {
        if (rails_mv_upload(B.device_handle(), 0, n_b, B_in, rows))
            return 200;
    }
    else
        B.random();

    const bool use_mass = params.get("Use mass matrix", false) ||
                          params.get("M-orthogonalization", false);
    CudaOperatorWrapper M = A;
    if (m_rowptr)
    {
        M = CudaOperatorWrapper(rows, global_rows, global_rows, row_offset, m_rowptr, m_colidx, m_vals);
        if (!M.device_handle())
            return 200;
    }
    else if (use_mass)
    {
        std::cerr << "rails_host_solve: \"Use mass matrix\" needs the CSR arrays of M" << std::endl;
        return 102;
    }

    CudaSolver solver(A, B, M);
    int ret = solver.set_parameters(params);
    if (ret)
        return 100 + ret;
    solver.set_verbose(std::getenv("RAILS_VERBOSE") != nullptr);
    if (n_null > 0 && N_in)
    {
        CudaMultiVectorWrapper N(global_rows, n_null);
        if (rails_mv_upload(N.device_handle(), 0, n_null, N_in, rows))
            return 200;
        solver.set_nullspace(N);
    }

    bool warm = params.get("Restart from solution", false);
    CudaMultiVectorWrapper V(global_rows, warm && k_io ? std::max(*k_io, 1) : 1);
    if (warm && k_io && V_io && *k_io > 0)
    {
        if (rails_mv_upload(V.device_handle(), 0, *k_io, V_io, rows))
            return 200;
    }
    HostDenseMatrix T;

    bool have_start = false, have_stop = false;
    long long launches_start = 0, launches_stop = 0;
    solver.set_observer([&](b200::IterationRecord const &rec) {
        if (rec.iteration == time_from)
        {
            have_start = rails_cuda_event_record(0) == 0;
            launches_start = rails_cuda_kernel_launches();
        }
        if (rec.iteration == time_to)
        {
            have_stop = rails_cuda_event_record(1) == 0;
            launches_stop = rails_cuda_kernel_launches();
        }
    });
    rails_cuda_sync();
    if (time_from == 0 && time_to > 0)
    {
        have_start = rails_cuda_event_record(0) == 0;
        launches_start = rails_cuda_kernel_launches();
    }
    auto t0 = std::chrono::steady_clock::now();
    if (setup_seconds)
        *setup_seconds = std::chrono::duration<double>(t0 - t_setup).count();

    ret = solver.solve(V, T);
    rails_cuda_sync();
    auto t1 = std::chrono::steady_clock::now();
    if (solve_seconds)
        *solve_seconds = std::chrono::duration<double>(t1 - t0).count();
    if (timed_ms)
    {
        *timed_ms = 0.0f;
        if (have_start && have_stop)
            rails_cuda_event_elapsed_ms(0, 1, timed_ms);
    }
    g_timed_launches = have_start && have_stop ? launches_stop - launches_start : 0;

    auto const &hist = solver.history();
    if (hist_len)
        *hist_len = (int)hist.size();
    for (int i = 0; i < std::min((int)hist.size(), hist_cap); ++i)
    {
        if (hist_abs) hist_abs[i] = hist[i].residual;
        if (hist_rel) hist_rel[i] = hist[i].relative;
        if (hist_k) hist_k[i] = hist[i].space_size;
        if (hist_sec) hist_sec[i] = hist[i].seconds;
    }
    std::ostringstream os;
    for (auto const &kv : solver.sections())
        os << kv.first << "=" << kv.second << ";";
    g_sections = os.str();
    if (dense_solve_seconds)
    {
        auto it = solver.sections().find("dense_solve");
        *dense_solve_seconds = it == solver.sections().end() ? 0.0 : it->second;
    }

    int k = V.N();
    if (k_io)
        *k_io = k;
    if (V_io && T_out && k <= v_cap_cols)
    {
        if (V.to_host(V_io, rows))
            return 200;
        for (int j = 0; j < k; ++j)
            for (int i = 0; i < k; ++i)
                T_out[i + (size_t)j * v_cap_cols] = T(i, j);
    }
    return ret;
}

int rails_host_solve_m(int rows, int global_rows, int row_offset,
                     const int *rowptr, const int *colidx, const double *vals,
                     const int *m_rowptr, const int *m_colidx, const double *m_vals,
                     int n_b, const double *B_in, unsigned seed,
                     int nparams, const char *const *keys, const double *pvals,
                     int hist_cap, double *hist_abs, double *hist_rel, int *hist_k,
                     double *hist_sec, int *hist_len,
                     int *k_io, int v_cap_cols, double *V_io, double *T_out,
                     int time_from, int time_to, float *timed_ms,
                     double *solve_seconds, double *setup_seconds, double *dense_solve_seconds)
{
    return rails_host_solve_x(rows, global_rows, row_offset, rowptr, colidx, vals, m_rowptr, m_colidx,
                              m_vals, n_b, B_in, seed, nparams, keys, pvals, hist_cap, hist_abs,
                              hist_rel, hist_k, hist_sec, hist_len, k_io, v_cap_cols, V_io, T_out,
                              time_from, time_to, timed_ms, solve_seconds, setup_seconds,
                              dense_solve_seconds, 0, nullptr);
}

int rails_host_solve(int rows, int global_rows, int row_offset,
                     const int *rowptr, const int *colidx, const double *vals,
                     int n_b, const double *B_in, unsigned seed,
                     int nparams, const char *const *keys, const double *pvals,
                     int hist_cap, double *hist_abs, double *hist_rel, int *hist_k,
                     double *hist_sec, int *hist_len,
                     int *k_io, int v_cap_cols, double *V_io, double *T_out,
                     int time_from, int time_to, float *timed_ms,
                     double *solve_seconds, double *setup_seconds, double *dense_solve_seconds)
{
    return rails_host_solve_m(rows, global_rows, row_offset, rowptr, colidx, vals, nullptr, nullptr,
                              nullptr, n_b, B_in, seed, nparams, keys, pvals, hist_cap, hist_abs,
                              hist_rel, hist_k, hist_sec, hist_len, k_io, v_cap_cols, V_io, T_out,
                              time_from, time_to, timed_ms, solve_seconds, setup_seconds,
                              dense_solve_seconds);
}

// "name=seconds;" list of the last solve's host-side section totals
const char *rails_host_last_sections(void) { return g_sections.c_str(); }

// cap the host BLAS pool as rails_host_solve does (idempotent); exported so
// that a caller driving the wrappers with its own loop sees the same k x k
// summation order
void rails_host_limit_blas_threads(void) { limit_blas_threads(); }

// kernels launched between the two timing marks of the last solve
long long rails_host_last_timed_launches(void) { return g_timed_launches; }

}
