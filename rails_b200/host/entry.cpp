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

#include <atomic>
#include <chrono>
#include <thread>
#include <dlfcn.h>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
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
long long g_timed_syncs = 0;

}

namespace
{

// A solve session: the operator(s), B and the solution factors stay RESIDENT on the
// device between solves, so that a continuation (new values on the same sparsity
// pattern, "Restart from solution" = 1: src/LyapunovSolver.hpp:108-121,
// test/LyapunovSolver_test.cpp:312-352, BASELINE config 5) neither re-uploads the
// operator nor round-trips V (m x k doubles: 17 GB at C5) through the host.
struct Session
{
    int rows = 0, global_rows = 0, row_offset = 0, n_b = 0;
    CudaOperatorWrapper A, M;
    bool has_M = false;
    CudaMultiVectorWrapper B, V, N;
    int n_null = 0;
    bool has_solution = false;
    HostDenseMatrix T;
    double setup_seconds = 0.0;
    std::atomic<int> *space_now = nullptr;      // updated every iteration when set (one-shot solve)
};

}

namespace
{
// set-up = everything between entry and the download that is not the solve
double s_setup_for_trace(std::chrono::steady_clock::time_point t0, std::chrono::steady_clock::time_point t1,
                         const double *solve_seconds)
{
    return std::chrono::duration<double>(t1 - t0).count() - (solve_seconds ? *solve_seconds : 0.0);
}
}

extern "C" {

// Uploads this rank's row block of A (and M), B (or draws it: B_in == NULL ->
// `std::srand(seed); B.random()`, the reference tests' way of making B,
// test/LyapunovSolver_test.cpp:66-71).  skip_draws: std::rand() values to discard
// after srand -- 1 when the caller made B itself with `srand(seed); B.random()`.
// Returns NULL on failure (rails_cuda_last_error has the reason).
void *rails_host_session_create(int rows, int global_rows, int row_offset,
                                const int *rowptr, const int *colidx, const double *vals,
                                const int *m_rowptr, const int *m_colidx, const double *m_vals,
                                int n_b, const double *B_in, unsigned seed, int skip_draws)
{
    if (!rails_cuda_is_initialized())
    {
        std::cerr << "rails_host_session_create: rails_cuda_init has not been called" << std::endl;
        return nullptr;
    }
    limit_blas_threads();
    auto t_setup = std::chrono::steady_clock::now();
    std::srand(seed);
    for (int i = 0; i < skip_draws; ++i)
        (void)std::rand();
    std::unique_ptr<Session> s(new Session());
    s->rows = rows; s->global_rows = global_rows; s->row_offset = row_offset; s->n_b = n_b;
    s->A = CudaOperatorWrapper(rows, global_rows, global_rows, row_offset, rowptr, colidx, vals);
    if (!s->A.device_handle())
        return nullptr;
    s->B = CudaMultiVectorWrapper(global_rows, n_b);
    if (B_in)
    {
        if (rails_mv_upload(s->B.device_handle(), 0, n_b, B_in, rows))
            return nullptr;
    }
    else
        s->B.random();
    s->M = s->A;
    if (m_rowptr)
    {
        s->M = CudaOperatorWrapper(rows, global_rows, global_rows, row_offset, m_rowptr, m_colidx, m_vals);
        if (!s->M.device_handle())
            return nullptr;
        s->has_M = true;
    }
    rails_cuda_sync();
    s->setup_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_setup).count();
    return s.release();
}

// New values of A on the SAME sparsity pattern, in the order of the arrays given at
// creation (a continuation step): row blocks, renumbered columns and halo plan stay.
int rails_host_session_update_operator(void *session, const double *vals)
{
    Session *s = static_cast<Session *>(session);
    if (!s || !vals)
        return 200;
    return rails_csr_update_values(s->A.device_handle(), vals) ? 200 : 0;
}

// this rank's rows of a space to project out of every vector that enters V
// (opts.nullspace of matlab/RAILSsolver.m:28-29); n_null = 0 clears it
int rails_host_session_set_nullspace(void *session, int n_null, const double *N_in)
{
    Session *s = static_cast<Session *>(session);
    if (!s)
        return 200;
    s->n_null = 0;
    if (n_null > 0 && N_in)
    {
        s->N = CudaMultiVectorWrapper(s->global_rows, n_null);
        if (rails_mv_upload(s->N.device_handle(), 0, n_null, N_in, s->rows))
            return 200;
        s->n_null = n_null;
    }
    return 0;
}

// a start space from host memory (k columns, this rank's rows, ld = rows) for the
// next solve with "Restart from solution" = 1; a session that has solved before
// already holds one on the device
int rails_host_session_set_start(void *session, int k, const double *V_in)
{
    Session *s = static_cast<Session *>(session);
    if (!s || k < 1 || !V_in)
        return 200;
    s->V = CudaMultiVectorWrapper(s->global_rows, k);
    if (rails_mv_upload(s->V.device_handle(), 0, k, V_in, s->rows))
        return 200;
    s->has_solution = true;
    return 0;
}

// One solve.  Returns the solver's code (0 converged, -1 stopped unconverged, 1 fell
// out of the loop), 100 + code for a rejected parameter list, 200 for a CUDA failure.
// With the parameter "Restart from solution" = 1 the solve starts from the basis the
// session holds (previous solve or rails_host_session_set_start), on the device.
//   timed region: CUDA events are recorded on the library stream when iteration
//       `time_from` and iteration `time_to` finish; their distance is returned in
//       *timed_ms (0 if not reached).
int rails_host_session_solve(void *session, int nparams, const char *const *keys, const double *pvals,
                             int hist_cap, double *hist_abs, double *hist_rel, int *hist_k,
                             double *hist_sec, int *hist_len, int *k_out,
                             int time_from, int time_to, float *timed_ms,
                             double *solve_seconds, double *dense_solve_seconds)
{
    Session *s = static_cast<Session *>(session);
    if (!s || !rails_cuda_is_initialized())
        return 200;
    KeyValueParameters params(nparams, keys, pvals);
    const bool use_mass = params.get("Use mass matrix", false) ||
                          params.get("M-orthogonalization", false);
    if (use_mass && !s->has_M)
    {
        std::cerr << "rails_host_solve: \"Use mass matrix\" needs the CSR arrays of M" << std::endl;
        return 102;
    }

    // "Random seed" >= 0 re-seeds std::rand() for this solve (start vector of V, one
    // Lanczos start vector per iteration); by default the stream simply continues, as
    // it does between two solves of a reference program
    const int reseed = params.get("Random seed", -1);
    if (reseed >= 0)
        std::srand((unsigned)reseed);

    CudaSolver solver(s->A, s->B, s->M);
    int ret = solver.set_parameters(params);
    if (ret)
        return 100 + ret;
    solver.set_verbose(std::getenv("RAILS_VERBOSE") != nullptr);
    if (s->n_null > 0)
        solver.set_nullspace(s->N);

    const bool warm = params.get("Restart from solution", false);
    if (!(warm && s->has_solution))
        s->V = CudaMultiVectorWrapper(s->global_rows, 1);

    bool have_start = false, have_stop = false;
    long long launches_start = 0, launches_stop = 0, syncs_start = 0, syncs_stop = 0;
    solver.set_observer([&](b200::IterationRecord const &rec) {
        if (s->space_now)
            s->space_now->store(rec.space_size, std::memory_order_relaxed);
        if (rec.iteration == time_from)
        {
            have_start = rails_cuda_event_record(0) == 0;
            launches_start = rails_cuda_kernel_launches();
            syncs_start = rails_cuda_host_syncs();
        }
        if (rec.iteration == time_to)
        {
            have_stop = rails_cuda_event_record(1) == 0;
            launches_stop = rails_cuda_kernel_launches();
            syncs_stop = rails_cuda_host_syncs();
        }
    });
    rails_cuda_sync();
    if (time_from == 0 && time_to > 0)
    {
        have_start = rails_cuda_event_record(0) == 0;
        launches_start = rails_cuda_kernel_launches();
        syncs_start = rails_cuda_host_syncs();
    }
    auto t0 = std::chrono::steady_clock::now();

    ret = solver.solve(s->V, s->T);
    rails_cuda_sync();
    s->has_solution = true;
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
    g_timed_syncs = have_start && have_stop ? syncs_stop - syncs_start : 0;

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
    if (k_out)
        *k_out = s->V.N();
    return ret;
}

// the factors of the last solution: V (this rank's rows, ld = rows) if it fits
// v_cap_cols columns, T as v_cap_cols x v_cap_cols.  Returns the space size, or
// -1 on failure / when the solution does not fit.
int rails_host_session_download(void *session, int v_cap_cols, double *V_out, double *T_out)
{
    Session *s = static_cast<Session *>(session);
    if (!s || !s->has_solution)
        return -1;
    const int k = s->V.N();
    if (!V_out || !T_out || k > v_cap_cols)
        return -1;
    if (s->V.to_host(V_out, s->rows))
        return -1;
    for (int j = 0; j < k; ++j)
        for (int i = 0; i < k; ++i)
            T_out[i + (size_t)j * v_cap_cols] = s->T(i, j);
    return k;
}

int rails_host_session_space_size(void *session)
{
    Session *s = static_cast<Session *>(session);
    return s && s->has_solution ? s->V.N() : -1;
}

double rails_host_session_setup_seconds(void *session)
{
    Session *s = static_cast<Session *>(session);
    return s ? s->setup_seconds : 0.0;
}

int rails_host_session_free(void *session)
{
    delete static_cast<Session *>(session);
    return 0;
}

// The one-shot form of the above: host CSR arrays and (optionally) a host B go in,
// the factors come back to host memory.
//
//   rows/global_rows/row_offset : this rank's row block of the square operator
//       (single rank: rows = global_rows, row_offset = 0); colidx is GLOBAL
//   B_in   : this rank's rows of B (rows x n_b, column-major, ld = rows), or NULL
//   V_io   : warm start (params "Restart from solution" = 1): on entry k_io
//            columns of V (this rank's rows, ld = rows); on exit the solution
//            basis if it fits v_cap_cols columns.  T_out is v_cap_cols^2.
//   m_rowptr/m_colidx/m_vals : this rank's rows of the mass matrix M, or NULL.  Used
//            only with "Use mass matrix" = 1 (A X M' + M X A' + B B' = 0,
//            matlab/RAILSsolver.m:368-392); otherwise M is stored and never read,
//            like the reference C++ loop.
//   n_null/N_in : null space to project out (rows x n_null, ld = rows), or 0/NULL
//   parameter "Skip random draws" = d: discard d std::rand() values after srand(seed)
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
    KeyValueParameters params(nparams, keys, pvals);
    const bool use_mass = params.get("Use mass matrix", false) ||
                          params.get("M-orthogonalization", false);
    if (use_mass && !m_rowptr)
    {
        std::cerr << "rails_host_solve: \"Use mass matrix\" needs the CSR arrays of M" << std::endl;
        return 102;
    }
    auto t_setup = std::chrono::steady_clock::now();
    const bool warm = params.get("Restart from solution", false);
    // The output block V_io is usually fresh memory: every page faults on its first
    // write, ~10 GB/s when that happens inside the download.  A helper thread touches
    // the columns the solution can occupy (the space size so far, plus two) while the
    // GPU works; the values are left as they are.
    std::atomic<int> space_now(0);
    std::atomic<bool> stop_touching(false);
    std::thread toucher;
    struct JoinToucher
    {
        std::atomic<bool> &stop; std::thread &t;
        ~JoinToucher() { stop.store(true); if (t.joinable()) t.join(); }
    } join_toucher{stop_touching, toucher};
    if (V_io && v_cap_cols > 0 && rows > 0 && !warm && !std::getenv("RAILS_NO_PREFAULT"))
        toucher = std::thread([&, rows, v_cap_cols, V_io] {
            const size_t col_bytes = (size_t)rows * sizeof(double);
            int done = 0;
            while (!stop_touching.load(std::memory_order_relaxed))
            {
                const int target = std::min(v_cap_cols, std::max(8, space_now.load(std::memory_order_relaxed) + 2));
                if (done >= target)
                {
                    std::this_thread::sleep_for(std::chrono::microseconds(200));
                    continue;
                }
                volatile char *col = reinterpret_cast<volatile char *>(V_io) + (size_t)done * col_bytes;
                for (size_t off = 0; off < col_bytes && !stop_touching.load(std::memory_order_relaxed); off += 4096)
                    col[off] = col[off];
                col[col_bytes - 1] = col[col_bytes - 1];
                ++done;
            }
        });
    Session *s = static_cast<Session *>(rails_host_session_create(
        rows, global_rows, row_offset, rowptr, colidx, vals, m_rowptr, m_colidx, m_vals, n_b, B_in, seed,
        params.get("Skip random draws", 0)));
    if (!s)
        return 200;
    std::unique_ptr<Session> owner(s);
    s->space_now = &space_now;
    if (n_null > 0 && N_in && rails_host_session_set_nullspace(s, n_null, N_in))
        return 200;
    if (warm && k_io && V_io && *k_io > 0 && rails_host_session_set_start(s, *k_io, V_io))
        return 200;
    rails_cuda_sync();
    if (setup_seconds)
        *setup_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_setup).count();

    int k = 0;
    int ret = rails_host_session_solve(s, nparams, keys, pvals, hist_cap, hist_abs, hist_rel, hist_k, hist_sec,
                                       hist_len, &k, time_from, time_to, timed_ms, solve_seconds,
                                       dense_solve_seconds);
    if (ret >= 100)
        return ret;
    if (k_io)
        *k_io = k;
    s->space_now = nullptr;
    stop_touching.store(true);
    if (toucher.joinable())
        toucher.join();
    auto t_down = std::chrono::steady_clock::now();
    if (V_io && T_out && k <= v_cap_cols && rails_host_session_download(s, v_cap_cols, V_io, T_out) < 0)
        return 200;
    // RAILS_E2E_TRACE=1: where the wall clock of a one-shot solve goes (stderr)
    if (std::getenv("RAILS_E2E_TRACE"))
    {
        auto now = std::chrono::steady_clock::now();
        owner.reset();
        auto freed = std::chrono::steady_clock::now();
        std::cerr << "rails_host_solve trace: set-up " << s_setup_for_trace(t_setup, t_down, solve_seconds)
                  << " s, solve " << (solve_seconds ? *solve_seconds : 0.0) << " s, download "
                  << std::chrono::duration<double>(now - t_down).count() << " s, free "
                  << std::chrono::duration<double>(freed - now).count() << " s" << std::endl;
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
// stream synchronisations of the library inside the same window
long long rails_host_last_timed_syncs(void) { return g_timed_syncs; }

}
