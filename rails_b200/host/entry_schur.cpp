// C entry points for problems with a singular mass matrix: the host sparse LU,
// the Schur-complement operator on the GPU and the solve through it
// (counterpart of the Schur path of /root/reference/src/main.cpp:75-116 and of
// matlab/RAILSschur.m + RAILSsolver.m as driven by matlab/test/test_MOC.m).
#include "CudaFusedLanczos.hpp"
#include "CudaMultiVectorWrapper.hpp"
#include "CudaOperatorWrapper.hpp"
#include "CudaSchurOperator.hpp"
#include "HostDenseMatrix.hpp"
#include "RailsSolver.hpp"
#include "SparseLU.hpp"

#include <chrono>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>

using namespace RAILS;

extern "C" void rails_host_limit_blas_threads(void);

namespace
{

class KeyValueParameters
{
    std::map<std::string, double> values_;

public:
    KeyValueParameters(int n, const char *const *keys, const double *vals)
    {
        for (int i = 0; i < n; ++i)
            values_[keys[i]] = vals[i];
    }
    bool has(std::string const &name) const { return values_.count(name) > 0; }
    void set(std::string const &name, double v) { values_[name] = v; }
    template<typename T>
    T get(std::string const &name, T def)
    {
        auto it = values_.find(name);
        return it == values_.end() ? def : (T)it->second;
    }
};

typedef b200::Solver<CudaOperatorWrapper, CudaMultiVectorWrapper, HostDenseMatrix> CudaSolver;

// [A border; border' 0] and M, B padded with zeros: the null space of A added as
// a border (SchurOperator's third constructor argument, src/SchurOperator.cpp:
// 25-36,158-175; add_border of matlab/test/test_MOC.m:137-160)
struct Bordered
{
    int n = 0;
    std::vector<int> rowptr, colidx;
    std::vector<double> vals, mdiag;
};

Bordered add_border(int n, const int *rowptr, const int *colidx, const double *vals,
                    const double *mdiag, int nb, const double *border)
{
    Bordered out;
    out.n = n + nb;
    out.rowptr.assign(1, 0);
    for (int i = 0; i < n; ++i)
    {
        for (int p = rowptr[i]; p < rowptr[i + 1]; ++p)
        {
            out.colidx.push_back(colidx[p]);
            out.vals.push_back(vals[p]);
        }
        for (int c = 0; c < nb; ++c)
            if (border[i + (size_t)c * n] != 0.0)
            {
                out.colidx.push_back(n + c);
                out.vals.push_back(border[i + (size_t)c * n]);
            }
        out.rowptr.push_back((int)out.colidx.size());
    }
    for (int c = 0; c < nb; ++c)
    {
        for (int i = 0; i < n; ++i)
            if (border[i + (size_t)c * n] != 0.0)
            {
                out.colidx.push_back(i);
                out.vals.push_back(border[i + (size_t)c * n]);
            }
        out.rowptr.push_back((int)out.colidx.size());
    }
    out.mdiag.assign(mdiag, mdiag + n);
    out.mdiag.resize(n + nb, 0.0);
    return out;
}

struct SchurObject
{
    CudaSchurOperator op;
    int n_total = 0;
};

}

extern "C" {

// ---- host sparse LU (no CUDA): factorisation checks on the CPU ------------------
void *rails_host_lu_factor(int n, const int *rowptr, const int *colidx, const double *vals,
                           double pivot_threshold, const int *order, int *info)
{
    SparseLU *lu = new SparseLU();
    int rc = lu->factor(n, rowptr, colidx, vals, pivot_threshold, order);
    if (info)
        *info = rc;
    return lu;
}

void rails_host_lu_free(void *h) { delete (SparseLU *)h; }

void rails_host_lu_solve(void *h, double *x, int nrhs, long long ld, int trans)
{
    ((SparseLU *)h)->solve(x, nrhs, ld, trans != 0);
}

// out: nnz(L) strict, nnz(U), off-diagonal pivots;  piv: min |pivot|, max |pivot|
void rails_host_lu_info(void *h, long long *out, double *piv)
{
    SparseLU *lu = (SparseLU *)h;
    out[0] = (long long)lu->Li.size();
    out[1] = (long long)lu->Ui.size();
    out[2] = lu->off_diagonal_pivots;
    if (piv)
    {
        piv[0] = lu->min_abs_pivot;
        piv[1] = lu->max_abs_pivot;
    }
}

void rails_host_lu_permutations(void *h, int *pinv, int *q)
{
    SparseLU *lu = (SparseLU *)h;
    std::memcpy(pinv, lu->pinv.data(), sizeof(int) * lu->n);
    std::memcpy(q, lu->q.data(), sizeof(int) * lu->n);
}

// level of every row in the forward sweep with L (level[i] = longest dependency
// chain ending in row i) and the number of stored entries of that row; what the
// device schedule of csrc/sptrsv.cu is built from (diagnostics, scripts/)
void rails_host_lu_row_levels(void *h, int *level, int *row_nnz)
{
    SparseLU *lu = (SparseLU *)h;
    const int n = lu->n;
    for (int i = 0; i < n; ++i)
    {
        level[i] = 0;
        row_nnz[i] = 0;
    }
    // L is stored by columns with rows > column: column j pushes its level to its rows
    for (int j = 0; j < n; ++j)
        for (int p = lu->Lp[j]; p < lu->Lp[j + 1]; ++p)
        {
            const int i = lu->Li[p];
            if (level[i] < level[j] + 1)
                level[i] = level[j] + 1;
            ++row_nnz[i];
        }
}

// the device schedule of these factors executed on the host (rails_lu_plan_emulate)
int rails_host_lu_plan_emulate(void *h, int trans, double *x, int nrhs, long long ld, long long *stats)
{
    SparseLU *lu = (SparseLU *)h;
    return rails_lu_plan_emulate(lu->n, lu->Lp.data(), lu->Li.data(), lu->Lx.data(), lu->Up.data(),
                                 lu->Ui.data(), lu->Ux.data(), lu->pinv.data(), lu->q.data(), trans, x,
                                 nrhs, ld, stats);
}

void rails_host_minimum_degree(int n, const int *rowptr, const int *colidx, int *order)
{
    std::vector<int> o = minimum_degree_order(n, rowptr, colidx);
    std::memcpy(order, o.data(), sizeof(int) * n);
}

void rails_host_nested_dissection(int n, const int *rowptr, const int *colidx, int leaf, int *order)
{
    std::vector<int> o = nested_dissection_order(n, rowptr, colidx, leaf);
    std::memcpy(order, o.data(), sizeof(int) * n);
}

// Solves with those factors ON THE GPU: x (n x nrhs host block, ld) is uploaded,
// solved `repeat` times from the same right-hand side (timed with CUDA events
// on the library stream, average returned in *ms_per_solve) and downloaded.
// stats: nnz(L+U), levels of L, U, U', L', launches per solve.
int rails_host_lu_device_solve(void *h, int trans, double *x, int nrhs, long long ld, int repeat,
                               float *ms_per_solve, long long *stats)
{
    if (!rails_cuda_is_initialized())
        return 200;
    SparseLU *lu = (SparseLU *)h;
    rails_lu_t F = nullptr;
    if (rails_lu_create(lu->n, lu->Lp.data(), lu->Li.data(), lu->Lx.data(), lu->Up.data(),
                        lu->Ui.data(), lu->Ux.data(), lu->pinv.data(), lu->q.data(), &F))
        return 200;
    CudaMultiVectorWrapper B(lu->n, nrhs), X(lu->n, nrhs);
    int rc = rails_mv_upload(B.device_handle(), 0, nrhs, x, ld);
    if (!rc)
        rc = rails_lu_solve(F, trans, B.device_handle(), 0, nrhs, X.device_handle(), 0);    // warm-up
    if (!rc && repeat > 0)
    {
        rails_cuda_event_record(2);
        for (int r = 0; r < repeat && !rc; ++r)
            rc = rails_lu_solve(F, trans, B.device_handle(), 0, nrhs, X.device_handle(), 0);
        rails_cuda_event_record(3);
        float ms = 0.0f;
        rails_cuda_event_elapsed_ms(2, 3, &ms);
        if (ms_per_solve)
            *ms_per_solve = ms / repeat;
    }
    if (!rc)
        rc = rails_mv_download(X.device_handle(), 0, nrhs, x, ld);
    if (stats)
    {
        stats[0] = rails_lu_nnz(F);
        for (int w = 0; w < 4; ++w)
            stats[1 + w] = rails_lu_levels(F, w);
        stats[5] = rails_lu_launches_per_solve(F, trans);
    }
    rails_lu_free(F);
    return rc ? 200 : 0;
}

// ---- the Schur-complement operator as an object (parity tests of its pieces) ------
void *rails_host_schur_create(int n, const int *rowptr, const int *colidx, const double *vals,
                              const double *mdiag, int n_border, const double *border,
                              double zero_tolerance, double scale, int *info)
{
    if (!rails_cuda_is_initialized())
    {
        if (info) *info = 200;
        return nullptr;
    }
    SchurObject *obj = new SchurObject();
    if (n_border > 0 && border)
    {
        Bordered b = add_border(n, rowptr, colidx, vals, mdiag, n_border, border);
        obj->op = CudaSchurOperator(b.n, b.rowptr.data(), b.colidx.data(), b.vals.data(),
                                    b.mdiag.data(), zero_tolerance, scale);
        obj->n_total = b.n;
    }
    else
    {
        obj->op = CudaSchurOperator(n, rowptr, colidx, vals, mdiag, zero_tolerance, scale);
        obj->n_total = n;
    }
    int rc = obj->op.Compute();
    if (info)
        *info = rc;
    return obj;
}

void rails_host_schur_free(void *h) { delete (SchurObject *)h; }

void rails_host_schur_sizes(void *h, int *n_total, int *n1, int *n2)
{
    SchurObject *obj = (SchurObject *)h;
    *n_total = obj->n_total;
    *n1 = obj->op.n1();
    *n2 = obj->op.n2();
}

void rails_host_schur_indices(void *h, int *idx1, int *idx2)
{
    SchurObject *obj = (SchurObject *)h;
    std::memcpy(idx1, obj->op.idx1().data(), sizeof(int) * obj->op.n1());
    std::memcpy(idx2, obj->op.idx2().data(), sizeof(int) * obj->op.n2());
}

// stats: nnz(L+U) of A11, levels of L, U, launches per A11 solve, off-diagonal
// pivots, operator applies (columns) so far
void rails_host_schur_stats(void *h, long long *stats, double *factor_seconds)
{
    SchurObject *obj = (SchurObject *)h;
    stats[0] = obj->op.factor_nnz();
    stats[1] = obj->op.factor_levels(0);
    stats[2] = obj->op.factor_levels(1);
    stats[3] = obj->op.launches_per_solve();
    stats[4] = obj->op.off_diagonal_pivots();
    stats[5] = obj->op.GetMVPs();
    if (factor_seconds)
        *factor_seconds = obj->op.factor_seconds();
}

// what = 0: Y = S X, 1: Y = S' X, 2: Y = S^-1 X  (n2 rows in and out);
//        3: Y = restrict(X) (n_total rows -> n2), 4: Y = prolongate(X) (n2 -> n_total).
// X, Y column-major host blocks with their row count as leading dimension.
// `repeat` > 0 additionally times that many further applies (CUDA events).
int rails_host_schur_apply(void *h, int what, const double *X, int c, double *Y, int repeat,
                           float *ms_per_apply)
{
    SchurObject *obj = (SchurObject *)h;
    const int rows_in = what == 3 ? obj->n_total : obj->op.n2();
    CudaMultiVectorWrapper x(rows_in, c);
    if (rails_mv_upload(x.device_handle(), 0, c, X, rows_in))
        return 200;
    auto run = [&]() {
        switch (what)
        {
        case 0: return obj->op.apply(x, false);
        case 1: return obj->op.apply(x, true);
        case 2: return obj->op.apply_inverse(x);
        case 3: return obj->op.restrict(x);
        default: return obj->op.prolongate(x);
        }
    };
    const long long errors0 = rails_cuda_error_count();
    CudaMultiVectorWrapper y = run();
    if (repeat > 0)
    {
        rails_cuda_event_record(2);
        for (int r = 0; r < repeat; ++r)
            CudaMultiVectorWrapper again = run();
        rails_cuda_event_record(3);
        float ms = 0.0f;
        rails_cuda_event_elapsed_ms(2, 3, &ms);
        if (ms_per_apply)
            *ms_per_apply = ms / repeat;
    }
    if (rails_cuda_error_count() != errors0)
        return 200;
    return y.to_host(Y, y.M()) ? 200 : 0;
}

// ---- the solve -----------------------------------------------------------------------
// A (n x n CSR), the diagonal of M (zeros mark the singular part), optionally a
// border (n x n_border, the null space of A), B (n x n_b) all on the host.
// Reduces to the Schur complement, runs the solver loop there with MS as mass
// matrix (switched on automatically unless MS = I) and returns
//     V_full : (n + n_border) x k, the prolongated basis: X = V_full T V_full'
//              solves the ORIGINAL equation (Vtrans, matlab/RAILSschur.m:68)
//     V_red  : n2 x k, the basis of the reduced problem (may be NULL)
//     T_out  : k x k with leading dimension v_cap_cols
//     k0, V0_full : warm start (parameter "Restart from solution" = 1): k0 columns
//              of a previous V_full ((n + n_border) rows); its rows idx2 are the
//              reduced basis (the prolongation leaves V itself there)
// Parameters as rails_host_solve plus "Projection method" (with the inverse
// S^-1 x = (A^-1 [0; x])(idx2) supplied, matlab/test/test_MOC.m:45-49).
// sizes: n1, n2;  stats as rails_host_schur_stats;  seconds: set-up, solve.
int rails_host_solve_schur(int n, const int *rowptr, const int *colidx, const double *vals,
                           const double *mdiag, int n_border, const double *border,
                           int n_b, const double *B_in, unsigned seed,
                           int nparams, const char *const *keys, const double *pvals,
                           int hist_cap, double *hist_abs, double *hist_rel, int *hist_k,
                           int *hist_len, int *k_out, int v_cap_cols, double *V_full,
                           double *V_red, double *T_out, int *sizes, long long *stats,
                           double *seconds, int k0, const double *V0_full)
{
    if (!rails_cuda_is_initialized())
    {
        std::cerr << "rails_host_solve_schur: rails_cuda_init has not been called" << std::endl;
        return 200;
    }
    rails_host_limit_blas_threads();
    auto t_setup = std::chrono::steady_clock::now();
    std::srand(seed);

    KeyValueParameters params(nparams, keys, pvals);
    int info = 0;
    SchurObject *obj = (SchurObject *)rails_host_schur_create(
        n, rowptr, colidx, vals, mdiag, n_border, border,
        params.get("Schur zero tolerance", 1e-12), params.get("Schur scale", 1.0), &info);
    if (!obj || info)
    {
        delete obj;
        return info ? 300 + info : 200;
    }
    CudaSchurOperator &schur = obj->op;
    const int n_total = obj->n_total;
    if (sizes)
    {
        sizes[0] = schur.n1();
        sizes[1] = schur.n2();
    }

    // B padded with zero rows for the border, then restricted
    CudaMultiVectorWrapper B(n_total, n_b);
    B = 0.0;
    {
        std::vector<double> padded((size_t)n_total * n_b, 0.0);
        for (int j = 0; j < n_b; ++j)
            std::memcpy(&padded[(size_t)j * n_total], B_in + (size_t)j * n, sizeof(double) * n);
        if (rails_mv_upload(B.device_handle(), 0, n_b, padded.data(), n_total))
        {
            delete obj;
            return 200;
        }
    }
    CudaMultiVectorWrapper BS = schur.restrict(B);

    if (!params.has("Use mass matrix") && !schur.mass_is_identity())
        params.set("Use mass matrix", 1.0);

    CudaOperatorWrapper S = schur.schur_operator();
    CudaSolver solver(S, BS, schur.mass_operator());
    int ret = solver.set_parameters(params);
    if (ret)
    {
        delete obj;
        return 100 + ret;
    }
    solver.set_verbose(std::getenv("RAILS_VERBOSE") != nullptr);
    if (params.get("Projection method", 1.0) > 1.0 + 1e-8)
        solver.set_inverse(schur.inverse_operator());

    CudaMultiVectorWrapper V(schur.n2(), 1);
    if (params.get("Restart from solution", false))
    {
        if (k0 < 1 || !V0_full)
        {
            std::cerr << "rails_host_solve_schur: \"Restart from solution\" needs a start basis" << std::endl;
            delete obj;
            return 102;
        }
        CudaMultiVectorWrapper full(n_total, k0);
        if (rails_mv_upload(full.device_handle(), 0, k0, V0_full, n_total))
        {
            delete obj;
            return 200;
        }
        V = schur.select_regular(full);
    }
    HostDenseMatrix T;
    rails_cuda_sync();
    auto t0 = std::chrono::steady_clock::now();
    ret = solver.solve(V, T);
    rails_cuda_sync();
    auto t1 = std::chrono::steady_clock::now();
    if (seconds)
    {
        seconds[0] = std::chrono::duration<double>(t0 - t_setup).count();
        seconds[1] = std::chrono::duration<double>(t1 - t0).count();
    }

    auto const &hist = solver.history();
    if (hist_len)
        *hist_len = (int)hist.size();
    for (int i = 0; i < std::min((int)hist.size(), hist_cap); ++i)
    {
        if (hist_abs) hist_abs[i] = hist[i].residual;
        if (hist_rel) hist_rel[i] = hist[i].relative;
        if (hist_k) hist_k[i] = hist[i].space_size;
    }

    const int k = V.N();
    if (k_out)
        *k_out = (T.M() >= k && T.N() >= k) ? k : 0;
    int rc = 0;
    if (k <= v_cap_cols && T.M() >= k && T.N() >= k)
    {
        if (V_red)
            rc |= V.to_host(V_red, schur.n2());
        if (V_full)
        {
            CudaMultiVectorWrapper Vf = schur.prolongate(V);
            rc |= Vf.to_host(V_full, n_total);
        }
        if (T_out)
            for (int j = 0; j < k; ++j)
                for (int i = 0; i < k; ++i)
                    T_out[i + (size_t)j * v_cap_cols] = T(i, j);
    }
    if (stats)
        rails_host_schur_stats(obj, stats, nullptr);
    delete obj;
    return rc ? 200 : ret;
}

}
