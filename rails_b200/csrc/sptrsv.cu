// Sparse LU solves on the device: the `A11 \ x` inside every apply of the
// Schur-complement operator (/root/reference/src/SchurOperator.cpp:204-208,
// Amesos_Klu::Solve; matlab/RAILSschur.m:32-33) and the `A \ x` of the inverse
// projection methods (matlab/RAILSschur.m:62-66).
//
// The factors P A Q = L U come from the host factorisation
// (rails_b200/host/SparseLU.cpp) once per operator; here they are kept in HBM
// as four row-oriented triangular matrices (L, U for A^-1; U', L' for A'^-1)
// with a level schedule each: row i is in level l when the longest dependency
// chain that ends in it has l rows.  A solve is
//     permute in -> sweep the levels of the first factor -> of the second ->
//     permute out,
// in place on an internal work block.  A level with many rows is one launch
// over all SMs (a warp per row, lanes over the row's entries, up to 4
// right-hand sides in registers); runs of consecutive small levels -- the long
// tail every elimination tree has near its root -- are walked by ONE CTA with a
// block barrier between levels, so the solve costs a handful of launches and
// not one per level.  Summation order inside a row is fixed (lane-strided
// partial sums, xor-shuffle tree): results are bit-reproducible.
//
// Bound: HBM/L2 latency for small factors, HBM bandwidth for large ones;
// algorithmic bytes per solve = 12 nnz(L + U) + 16 n c (c right-hand sides).
#include "internal.h"

#include <algorithm>

using rails::ctx;

namespace
{

constexpr int LEVEL_THREADS = 256;         // multi-CTA level kernel: 8 warps
constexpr int CHAIN_THREADS = 1024;        // single-CTA chain kernel: 32 warps
constexpr int SMALL_LEVEL_ROWS = 128;      // levels up to this size go into a chain

struct Triangle
{
    int n = 0;
    long long nnz = 0;
    int *rowptr = nullptr;
    int *col = nullptr;
    double *val = nullptr;
    double *diag = nullptr;                // nullptr: unit diagonal
    int *order = nullptr;                  // rows sorted by level
    int *lvlptr = nullptr;                 // device copy of the level offsets
    int nlevels = 0;
    struct Segment { int lvl0, lvl1, row0, row1; bool chain; };
    std::vector<Segment> segments;
    size_t bytes = 0;
};

template<int CT>
__device__ __forceinline__ void solve_row(const int *__restrict__ rowptr, const int *__restrict__ col,
                                          const double *__restrict__ val,
                                          const double *__restrict__ diag, int row, double *X,
                                          long long ld, int c0, int lane)
{
    double acc[CT];
#pragma unroll
    for (int c = 0; c < CT; ++c)
        acc[c] = 0.0;
    const int pend = rowptr[row + 1];
    for (int p = rowptr[row] + lane; p < pend; p += 32)
    {
        const int j = col[p];
        const double v = val[p];
#pragma unroll
        for (int c = 0; c < CT; ++c)
            acc[c] = fma(v, __ldcg(X + j + (size_t)(c0 + c) * ld), acc[c]);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1)
#pragma unroll
        for (int c = 0; c < CT; ++c)
            acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], off);
    if (lane == 0)
    {
        const double d = diag ? diag[row] : 1.0;
#pragma unroll
        for (int c = 0; c < CT; ++c)
        {
            double *x = X + row + (size_t)(c0 + c) * ld;
            const double r = __ldcg(x) - acc[c];
            *x = diag ? r / d : r;
        }
    }
}

__device__ __forceinline__ void solve_row_all(const int *rowptr, const int *col, const double *val,
                                              const double *diag, int row, double *X, long long ld,
                                              int nc, int lane)
{
    int c0 = 0;
    for (; c0 + 4 <= nc; c0 += 4)
        solve_row<4>(rowptr, col, val, diag, row, X, ld, c0, lane);
    if (c0 + 2 <= nc)
    {
        solve_row<2>(rowptr, col, val, diag, row, X, ld, c0, lane);
        c0 += 2;
    }
    if (c0 < nc)
        solve_row<1>(rowptr, col, val, diag, row, X, ld, c0, lane);
}

// one level, rows order[r0 .. r1): a warp per row, grid-strided
__global__ void __launch_bounds__(LEVEL_THREADS)
sptrsv_level_kernel(const int *__restrict__ rowptr, const int *__restrict__ col,
                    const double *__restrict__ val, const double *__restrict__ diag,
                    const int *__restrict__ order, int r0, int r1, double *X, long long ld, int nc)
{
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * LEVEL_THREADS + threadIdx.x) >> 5;
    const int nwarps = (gridDim.x * LEVEL_THREADS) >> 5;
    for (int r = r0 + warp; r < r1; r += nwarps)
        solve_row_all(rowptr, col, val, diag, order[r], X, ld, nc, lane);
}

// levels [l0, l1) walked by one CTA, a block barrier between levels
__global__ void __launch_bounds__(CHAIN_THREADS)
sptrsv_chain_kernel(const int *__restrict__ rowptr, const int *__restrict__ col,
                    const double *__restrict__ val, const double *__restrict__ diag,
                    const int *__restrict__ order, const int *__restrict__ lvlptr, int l0, int l1,
                    double *X, long long ld, int nc)
{
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    constexpr int nwarps = CHAIN_THREADS / 32;
    for (int l = l0; l < l1; ++l)
    {
        const int r1 = lvlptr[l + 1];
        for (int r = lvlptr[l] + warp; r < r1; r += nwarps)
            solve_row_all(rowptr, col, val, diag, order[r], X, ld, nc, lane);
        __syncthreads();
    }
}

// dst[map[i]] = src[i] (scatter) or dst[i] = src[map[i]] (gather), column by column
__global__ void __launch_bounds__(256)
permute_rows_kernel(double *__restrict__ dst, long long ldd, const double *__restrict__ src,
                    long long lds, const int *__restrict__ map, long long n, int scatter,
                    double scale)
{
    double *d = dst + (size_t)blockIdx.y * ldd;
    const double *s = src + (size_t)blockIdx.y * lds;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256)
    {
        if (scatter)
            d[map[i]] = scale * s[i];
        else
            d[i] = scale * s[map[i]];
    }
}

int upload(const void *host, size_t bytes, void **dev)
{
    RAILS_CUDA_TRY(cudaMalloc(dev, bytes ? bytes : 8));
    if (bytes)
        RAILS_CUDA_TRY(cudaMemcpyAsync(*dev, host, bytes, cudaMemcpyHostToDevice, ctx().stream));
    return RAILS_OK;
}

// Builds the device triangle from a host CSR-like triple holding the STRICT
// part (entries of row i have col < i when `lower`, col > i otherwise).
int build_triangle(int n, const std::vector<int> &ptr, const std::vector<int> &idx,
                   const std::vector<double> &val, const std::vector<double> *diag, bool lower,
                   Triangle &T)
{
    T.n = n;
    T.nnz = (long long)idx.size();
    std::vector<int> level(n, 0);
    int nlevels = n > 0 ? 1 : 0;
    for (int s = 0; s < n; ++s)
    {
        const int i = lower ? s : n - 1 - s;
        int l = 0;
        for (int p = ptr[i]; p < ptr[i + 1]; ++p)
            l = std::max(l, level[idx[p]] + 1);
        level[i] = l;
        nlevels = std::max(nlevels, l + 1);
    }
    std::vector<int> lvlptr(nlevels + 1, 0), order(n);
    for (int i = 0; i < n; ++i)
        ++lvlptr[level[i] + 1];
    for (int l = 0; l < nlevels; ++l)
        lvlptr[l + 1] += lvlptr[l];
    {
        std::vector<int> next(lvlptr.begin(), lvlptr.end() - 1);
        for (int i = 0; i < n; ++i)
            order[next[level[i]]++] = i;
    }
    T.nlevels = nlevels;
    T.segments.clear();
    for (int l = 0; l < nlevels;)
    {
        const int rows = lvlptr[l + 1] - lvlptr[l];
        Triangle::Segment seg;
        seg.lvl0 = l;
        if (rows > SMALL_LEVEL_ROWS)
        {
            seg.lvl1 = l + 1;
            seg.chain = false;
        }
        else
        {
            int e = l + 1;
            while (e < nlevels && lvlptr[e + 1] - lvlptr[e] <= SMALL_LEVEL_ROWS)
                ++e;
            seg.lvl1 = e;
            seg.chain = true;
        }
        seg.row0 = lvlptr[seg.lvl0];
        seg.row1 = lvlptr[seg.lvl1];
        T.segments.push_back(seg);
        l = seg.lvl1;
    }

    int rc;
    if ((rc = upload(ptr.data(), (size_t)(n + 1) * sizeof(int), (void **)&T.rowptr))) return rc;
    if ((rc = upload(idx.data(), idx.size() * sizeof(int), (void **)&T.col))) return rc;
    if ((rc = upload(val.data(), val.size() * sizeof(double), (void **)&T.val))) return rc;
    if (diag && (rc = upload(diag->data(), (size_t)n * sizeof(double), (void **)&T.diag))) return rc;
    if ((rc = upload(order.data(), (size_t)n * sizeof(int), (void **)&T.order))) return rc;
    if ((rc = upload(lvlptr.data(), lvlptr.size() * sizeof(int), (void **)&T.lvlptr))) return rc;
    RAILS_CUDA_TRY(cudaStreamSynchronize(ctx().stream));      // host vectors die with the caller
    T.bytes = (size_t)(n + 1) * 4 + idx.size() * 12 + (diag ? (size_t)n * 8 : 0) + (size_t)n * 4 +
              lvlptr.size() * 4;
    return RAILS_OK;
}

void free_triangle(Triangle &T)
{
    cudaFree(T.rowptr); cudaFree(T.col); cudaFree(T.val); cudaFree(T.diag);
    cudaFree(T.order); cudaFree(T.lvlptr);
    T = Triangle();
}

int sweep(const Triangle &T, double *X, long long ld, int nc)
{
    for (size_t s = 0; s < T.segments.size(); ++s)
    {
        const Triangle::Segment &seg = T.segments[s];
        if (seg.chain)
            RAILS_LAUNCH(sptrsv_chain_kernel, 1, CHAIN_THREADS, 0, T.rowptr, T.col, T.val, T.diag,
                         T.order, T.lvlptr, seg.lvl0, seg.lvl1, X, ld, nc);
        else
        {
            const int rows = seg.row1 - seg.row0;
            const int warps_per_cta = LEVEL_THREADS / 32;
            int grid = (rows + warps_per_cta - 1) / warps_per_cta;
            grid = std::min(grid, ctx().sm_count * 8);
            RAILS_LAUNCH(sptrsv_level_kernel, grid, LEVEL_THREADS, 0, T.rowptr, T.col, T.val, T.diag,
                         T.order, seg.row0, seg.row1, X, ld, nc);
        }
    }
    RAILS_CHECK_LAUNCH();
    return RAILS_OK;
}

// CSC (colptr, rowidx, vals) -> CSR, optionally pulling the diagonal out
void transpose_strict(int n, const int *cp, const int *ri, const double *cv, std::vector<int> &ptr,
                      std::vector<int> &idx, std::vector<double> &val, std::vector<double> *diag)
{
    ptr.assign(n + 1, 0);
    for (int j = 0; j < n; ++j)
        for (int p = cp[j]; p < cp[j + 1]; ++p)
            if (ri[p] != j)
                ++ptr[ri[p] + 1];
    for (int i = 0; i < n; ++i)
        ptr[i + 1] += ptr[i];
    idx.resize(ptr[n]);
    val.resize(ptr[n]);
    std::vector<int> next(ptr.begin(), ptr.end() - 1);
    for (int j = 0; j < n; ++j)
        for (int p = cp[j]; p < cp[j + 1]; ++p)
        {
            if (ri[p] == j)
            {
                if (diag) (*diag)[j] = cv[p];
                continue;
            }
            const int d = next[ri[p]]++;
            idx[d] = j;
            val[d] = cv[p];
        }
}

void copy_strict(int n, const int *cp, const int *ri, const double *cv, std::vector<int> &ptr,
                 std::vector<int> &idx, std::vector<double> &val)
{
    ptr.assign(n + 1, 0);
    idx.clear();
    val.clear();
    for (int j = 0; j < n; ++j)
    {
        for (int p = cp[j]; p < cp[j + 1]; ++p)
            if (ri[p] != j)
            {
                idx.push_back(ri[p]);
                val.push_back(cv[p]);
            }
        ptr[j + 1] = (int)idx.size();
    }
}

}

struct rails_lu_s
{
    int n = 0;
    Triangle L, U, Ut, Lt;
    int *pinv = nullptr;
    int *q = nullptr;
    double *work = nullptr;
    int work_cols = 0;
    long long ld = 0;
    size_t bytes = 0;
};

struct rails_idx_s
{
    long long n = 0;
    long long max_index = -1;
    int *idx = nullptr;
};

extern "C" int rails_lu_create(int n, const int *Lp, const int *Li, const double *Lx,
                               const int *Up, const int *Ui, const double *Ux,
                               const int *pinv, const int *q, rails_lu_t *out)
{
    RAILS_REQUIRE_READY();
    if (n < 0 || !Lp || !Up || !pinv || !q || !out)
        return rails::fail(RAILS_ERR_ARG, "rails_lu_create: null argument or n = %d", n);
    // structure checks: L strictly lower, U upper with a non-zero diagonal
    std::vector<double> diag(n, 0.0);
    for (int j = 0; j < n; ++j)
    {
        for (int p = Lp[j]; p < Lp[j + 1]; ++p)
            if (Li[p] <= j || Li[p] >= n)
                return rails::fail(RAILS_ERR_ARG, "rails_lu_create: L(%d, %d) is not below the diagonal", Li[p], j);
        bool have = false;
        for (int p = Up[j]; p < Up[j + 1]; ++p)
        {
            if (Ui[p] > j || Ui[p] < 0)
                return rails::fail(RAILS_ERR_ARG, "rails_lu_create: U(%d, %d) is below the diagonal", Ui[p], j);
            if (Ui[p] == j && Ux[p] != 0.0)
                have = true;
        }
        if (!have)
            return rails::fail(RAILS_ERR_ARG, "rails_lu_create: zero pivot in column %d", j);
        if (pinv[j] < 0 || pinv[j] >= n || q[j] < 0 || q[j] >= n)
            return rails::fail(RAILS_ERR_ARG, "rails_lu_create: bad permutation entry %d", j);
    }

    rails_lu_s *F = new rails_lu_s();
    F->n = n;
    std::vector<int> ptr, idx;
    std::vector<double> val;
    int rc = RAILS_OK;
    do
    {
        // L y = b: rows of L (forward)
        transpose_strict(n, Lp, Li, Lx, ptr, idx, val, nullptr);
        if ((rc = build_triangle(n, ptr, idx, val, nullptr, true, F->L))) break;
        // U x = y: rows of U (backward)
        transpose_strict(n, Up, Ui, Ux, ptr, idx, val, &diag);
        if ((rc = build_triangle(n, ptr, idx, val, &diag, false, F->U))) break;
        // U' w = z: rows of U' are the columns of U (forward)
        copy_strict(n, Up, Ui, Ux, ptr, idx, val);
        if ((rc = build_triangle(n, ptr, idx, val, &diag, true, F->Ut))) break;
        // L' v = w: rows of L' are the columns of L (backward)
        copy_strict(n, Lp, Li, Lx, ptr, idx, val);
        if ((rc = build_triangle(n, ptr, idx, val, nullptr, false, F->Lt))) break;
        if ((rc = upload(pinv, (size_t)n * sizeof(int), (void **)&F->pinv))) break;
        if ((rc = upload(q, (size_t)n * sizeof(int), (void **)&F->q))) break;
        cudaError_t e = cudaStreamSynchronize(ctx().stream);
        if (e != cudaSuccess)
            rc = rails::fail(RAILS_ERR_CUDA, "rails_lu_create: %s", cudaGetErrorString(e));
    } while (false);
    if (rc)
    {
        rails_lu_free(F);
        return rc;
    }
    F->ld = (((long long)n + 15) / 16) * 16;
    if (F->ld == 0) F->ld = 16;
    F->bytes = F->L.bytes + F->U.bytes + F->Ut.bytes + F->Lt.bytes + (size_t)n * 8;
    ctx().bytes_in_use.fetch_add((long long)F->bytes);
    *out = F;
    return RAILS_OK;
}

extern "C" int rails_lu_free(rails_lu_t F)
{
    if (!F)
        return RAILS_OK;
    if (ctx().ready)
    {
        cudaStreamSynchronize(ctx().stream);
        free_triangle(F->L); free_triangle(F->U); free_triangle(F->Ut); free_triangle(F->Lt);
        cudaFree(F->pinv); cudaFree(F->q); cudaFree(F->work);
        ctx().bytes_in_use.fetch_sub((long long)F->bytes);
    }
    delete F;
    return RAILS_OK;
}

extern "C" long long rails_lu_nnz(rails_lu_t F)
{
    return F ? F->L.nnz + F->U.nnz + F->n : 0;
}

extern "C" int rails_lu_levels(rails_lu_t F, int which)
{
    if (!F) return 0;
    const Triangle *T = which == 0 ? &F->L : which == 1 ? &F->U : which == 2 ? &F->Ut : &F->Lt;
    return T->nlevels;
}

extern "C" int rails_lu_launches_per_solve(rails_lu_t F, int trans)
{
    if (!F) return 0;
    return 2 + (int)(trans ? F->Ut.segments.size() + F->Lt.segments.size()
                           : F->L.segments.size() + F->U.segments.size());
}

extern "C" int rails_lu_solve(rails_lu_t F, int trans, rails_mv_t B, int bc0, int nc,
                              rails_mv_t X, int xc0)
{
    RAILS_REQUIRE_READY();
    if (!F || !B || !X || nc < 0 || bc0 < 0 || xc0 < 0 || bc0 + nc > B->cols || xc0 + nc > X->cols)
        return rails::fail(RAILS_ERR_ARG, "rails_lu_solve: bad handle or column range");
    if (B->rows != F->n || X->rows != F->n)
        return rails::fail(RAILS_ERR_ARG, "rails_lu_solve: factors are %d x %d, B has %lld rows, X %lld",
                           F->n, F->n, B->rows, X->rows);
    if (nc == 0 || F->n == 0)
        return RAILS_OK;
    if (nc > F->work_cols)
    {
        RAILS_CUDA_TRY(cudaStreamSynchronize(ctx().stream));
        cudaFree(F->work);
        F->work = nullptr;
        F->work_cols = 0;
        RAILS_CUDA_TRY(cudaMalloc((void **)&F->work, (size_t)F->ld * nc * sizeof(double)));
        F->work_cols = nc;
    }
    const long long n = F->n;
    const dim3 pgrid((unsigned)std::min<long long>((n + 255) / 256, 4LL * ctx().sm_count), (unsigned)nc);
    double *W = F->work;
    const double *b = B->data + (size_t)bc0 * B->ld;
    double *x = X->data + (size_t)xc0 * X->ld;
    int rc;
    if (!trans)
    {
        // A x = b:  w[pinv[i]] = b[i];  L, then U;  x[q[j]] = w[j]
        RAILS_LAUNCH(permute_rows_kernel, pgrid, 256, 0, W, F->ld, b, B->ld, F->pinv, n, 1, 1.0);
        if ((rc = sweep(F->L, W, F->ld, nc))) return rc;
        if ((rc = sweep(F->U, W, F->ld, nc))) return rc;
        RAILS_LAUNCH(permute_rows_kernel, pgrid, 256, 0, x, X->ld, W, F->ld, F->q, n, 1, 1.0);
    }
    else
    {
        // A' x = b:  w[j] = b[q[j]];  U', then L';  x[i] = w[pinv[i]]
        RAILS_LAUNCH(permute_rows_kernel, pgrid, 256, 0, W, F->ld, b, B->ld, F->q, n, 0, 1.0);
        if ((rc = sweep(F->Ut, W, F->ld, nc))) return rc;
        if ((rc = sweep(F->Lt, W, F->ld, nc))) return rc;
        RAILS_LAUNCH(permute_rows_kernel, pgrid, 256, 0, x, X->ld, W, F->ld, F->pinv, n, 0, 1.0);
    }
    RAILS_CHECK_LAUNCH();
    return RAILS_OK;
}

extern "C" int rails_idx_create(long long n, const int *host_idx, rails_idx_t *out)
{
    RAILS_REQUIRE_READY();
    if (n < 0 || (n > 0 && !host_idx) || !out)
        return rails::fail(RAILS_ERR_ARG, "rails_idx_create(%lld)", n);
    rails_idx_s *I = new rails_idx_s();
    I->n = n;
    for (long long i = 0; i < n; ++i)
    {
        if (host_idx[i] < 0)
        {
            delete I;
            return rails::fail(RAILS_ERR_ARG, "rails_idx_create: negative index at %lld", i);
        }
        I->max_index = std::max<long long>(I->max_index, host_idx[i]);
    }
    int rc = upload(host_idx, (size_t)n * sizeof(int), (void **)&I->idx);
    if (!rc)
    {
        cudaError_t e = cudaStreamSynchronize(ctx().stream);
        if (e != cudaSuccess)
            rc = rails::fail(RAILS_ERR_CUDA, "rails_idx_create: %s", cudaGetErrorString(e));
    }
    if (rc)
    {
        cudaFree(I->idx);
        delete I;
        return rc;
    }
    *out = I;
    return RAILS_OK;
}

extern "C" int rails_idx_free(rails_idx_t I)
{
    if (!I)
        return RAILS_OK;
    if (ctx().ready)
    {
        cudaStreamSynchronize(ctx().stream);
        cudaFree(I->idx);
    }
    delete I;
    return RAILS_OK;
}

extern "C" int rails_mv_gather_rows(rails_mv_t dst, int dc0, rails_mv_t src, int sc0, int nc,
                                    rails_idx_t I, double scale)
{
    RAILS_REQUIRE_READY();
    if (!dst || !src || !I || nc < 0 || dc0 < 0 || sc0 < 0 || dc0 + nc > dst->cols || sc0 + nc > src->cols)
        return rails::fail(RAILS_ERR_ARG, "rails_mv_gather_rows: bad handle or column range");
    if (dst->rows != I->n || I->max_index >= src->rows)
        return rails::fail(RAILS_ERR_ARG, "rails_mv_gather_rows: %lld indices (max %lld) from %lld rows into %lld",
                           I->n, I->max_index, src->rows, dst->rows);
    if (nc == 0 || I->n == 0)
        return RAILS_OK;
    const dim3 grid((unsigned)std::min<long long>((I->n + 255) / 256, 4LL * ctx().sm_count), (unsigned)nc);
    RAILS_LAUNCH(permute_rows_kernel, grid, 256, 0, dst->data + (size_t)dc0 * dst->ld, dst->ld,
                 src->data + (size_t)sc0 * src->ld, src->ld, I->idx, I->n, 0, scale);
    RAILS_CHECK_LAUNCH();
    return RAILS_OK;
}

extern "C" int rails_mv_scatter_rows(rails_mv_t dst, int dc0, rails_mv_t src, int sc0, int nc,
                                     rails_idx_t I, double scale)
{
    RAILS_REQUIRE_READY();
    if (!dst || !src || !I || nc < 0 || dc0 < 0 || sc0 < 0 || dc0 + nc > dst->cols || sc0 + nc > src->cols)
        return rails::fail(RAILS_ERR_ARG, "rails_mv_scatter_rows: bad handle or column range");
    if (src->rows != I->n || I->max_index >= dst->rows)
        return rails::fail(RAILS_ERR_ARG, "rails_mv_scatter_rows: %lld indices (max %lld) from %lld rows into %lld",
                           I->n, I->max_index, src->rows, dst->rows);
    if (nc == 0 || I->n == 0)
        return RAILS_OK;
    const dim3 grid((unsigned)std::min<long long>((I->n + 255) / 256, 4LL * ctx().sm_count), (unsigned)nc);
    RAILS_LAUNCH(permute_rows_kernel, grid, 256, 0, dst->data + (size_t)dc0 * dst->ld, dst->ld,
                 src->data + (size_t)sc0 * src->ld, src->ld, I->idx, I->n, 1, scale);
    RAILS_CHECK_LAUNCH();
    return RAILS_OK;
}
