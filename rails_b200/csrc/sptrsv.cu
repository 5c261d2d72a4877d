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

#include <cuda_pipeline.h>

#include <algorithm>

using rails::ctx;

namespace
{

constexpr int LEVEL_THREADS = 256;         // multi-CTA level kernel: 8 warps
constexpr int CHAIN_THREADS = 512;         // single-CTA chain kernel: 16 warps, 128 registers each
constexpr int SMALL_LEVEL_ROWS = 128;      // levels up to this size go into a chain
// a chain is cut into windows whose matrix data fit one shared-memory stage
constexpr int WIN_ENTRIES = 8192;          // col (4 B) + val (8 B) per entry: 96 KB per stage
constexpr int WIN_ROWS = 256;              // int4 descriptors: 4 KB
constexpr int WIN_LEVELS = 64;

struct Triangle
{
    int n = 0;
    long long nnz = 0;
    // rows in level order: rowinfo[r] = {row, first entry, end, 0}; col / val
    // hold the strict part of the rows in that same order, so a sweep streams
    // through them front to back
    int4 *rowinfo = nullptr;
    int *col = nullptr;
    double *val = nullptr;
    double *diag = nullptr;                // by row; nullptr: unit diagonal
    int *lvlptr = nullptr;                 // device copy of the level offsets
    int *winptr = nullptr;                 // chain windows: first level of window w; bit 30 set when
                                           // the window is one level too large to stage
    int nlevels = 0;
    struct Segment { int lvl0, lvl1, row0, row1; bool chain; int win0, win1; };
    std::vector<Segment> segments;
    size_t bytes = 0;
};

template<int CT>
__device__ __forceinline__ void solve_row(const int4 info, const int *__restrict__ col,
                                          const double *__restrict__ val,
                                          const double *__restrict__ diag, double *X,
                                          long long ld, int c0, int lane)
{
    double acc[CT];
#pragma unroll
    for (int c = 0; c < CT; ++c)
        acc[c] = 0.0;
    const int row = info.x;
    for (int p = info.y + lane; p < info.z; p += 32)
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

__device__ __forceinline__ void solve_row_all(const int4 info, const int *col, const double *val,
                                              const double *diag, double *X, long long ld, int nc,
                                              int lane)
{
    int c0 = 0;
    for (; c0 + 4 <= nc; c0 += 4)
        solve_row<4>(info, col, val, diag, X, ld, c0, lane);
    if (c0 + 2 <= nc)
    {
        solve_row<2>(info, col, val, diag, X, ld, c0, lane);
        c0 += 2;
    }
    if (c0 < nc)
        solve_row<1>(info, col, val, diag, X, ld, c0, lane);
}

// one level, rows r0 .. r1 (level order): a warp per row, grid-strided
__global__ void __launch_bounds__(LEVEL_THREADS)
sptrsv_level_kernel(const int4 *__restrict__ rowinfo, const int *__restrict__ col,
                    const double *__restrict__ val, const double *__restrict__ diag, int r0, int r1,
                    double *X, long long ld, int nc)
{
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * LEVEL_THREADS + threadIdx.x) >> 5;
    const int nwarps = (gridDim.x * LEVEL_THREADS) >> 5;
    for (int r = r0 + warp; r < r1; r += nwarps)
        solve_row_all(rowinfo[r], col, val, diag, X, ld, nc, lane);
}

// A run of consecutive small levels walked by one CTA, a block barrier between
// levels.
//
// These are the levels near the root of the elimination tree: few rows (often
// one), each with hundreds of entries, every level waiting for the one before.
// What such a level costs is a chain of memory latencies, so everything that
// does not depend on the previous level is taken off that chain:
//   * the run is cut (on the host) into windows of up to WIN_LEVELS levels whose
//     row descriptors and matrix entries fit one shared-memory stage; they are
//     contiguous in memory because the factor is stored in level order.  While
//     the CTA works on one window, cp.async copies the next into the other
//     stage.  Inside a window only the gather of X, a warp reduction and two
//     barriers stand between consecutive levels;
//   * a level with fewer than 8 rows is worked on by the whole CTA: each row
//     gets a team of 16 / pow2(rows) warps that split its entries, the per-warp
//     partial sums are combined in fixed order through shared memory;
//   * a level too large for a stage is a window of its own and is read from
//     global memory, a warp per row.
struct WindowStage
{
    int lvl[WIN_LEVELS + 1];     // row offsets of the window's levels
    int4 info[WIN_ROWS];
    int col[WIN_ENTRIES];
    double val[WIN_ENTRIES];
};

struct WindowShape
{
    int la, lb;        // levels
    int ra, rb;        // rows (level order)
    int ea, eb;        // entries
    bool direct;
};

constexpr int WIN_DIRECT = 1 << 30;

__device__ __forceinline__ WindowShape window_shape(const int4 *__restrict__ rowinfo,
                                                    const int *__restrict__ lvlptr,
                                                    const int *__restrict__ winptr, int w)
{
    WindowShape s;
    const int a = winptr[w];
    s.direct = (a & WIN_DIRECT) != 0;
    s.la = a & ~WIN_DIRECT;
    s.lb = winptr[w + 1] & ~WIN_DIRECT;
    s.ra = lvlptr[s.la];
    s.rb = lvlptr[s.lb];
    s.ea = rowinfo[s.ra].y;
    s.eb = rowinfo[s.rb - 1].z;
    return s;
}

__device__ __forceinline__ void window_fetch(WindowStage &st, const WindowShape &s,
                                             const int4 *__restrict__ rowinfo,
                                             const int *__restrict__ col,
                                             const double *__restrict__ val,
                                             const int *__restrict__ lvlptr)
{
    if (s.direct)
        return;
    for (int i = threadIdx.x; i <= s.lb - s.la; i += CHAIN_THREADS)
        __pipeline_memcpy_async(&st.lvl[i], lvlptr + s.la + i, 4);
    for (int i = threadIdx.x; i < s.rb - s.ra; i += CHAIN_THREADS)
        __pipeline_memcpy_async(&st.info[i], rowinfo + s.ra + i, 16);
    for (int i = threadIdx.x; i < s.eb - s.ea; i += CHAIN_THREADS)
    {
        __pipeline_memcpy_async(&st.col[i], col + s.ea + i, 4);
        __pipeline_memcpy_async(&st.val[i], val + s.ea + i, 8);
    }
}

__global__ void __launch_bounds__(CHAIN_THREADS, 1)
sptrsv_chain_kernel(const int4 *__restrict__ rowinfo, const int *__restrict__ col,
                    const double *__restrict__ val, const double *__restrict__ diag,
                    const int *__restrict__ lvlptr, const int *__restrict__ winptr, int w0, int w1,
                    double *X, long long ld, int nc)
{
    extern __shared__ __align__(16) unsigned char chain_smem[];
    WindowStage *stages = reinterpret_cast<WindowStage *>(chain_smem);
    __shared__ double partial[CHAIN_THREADS / 32][4];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    constexpr int nwarps = CHAIN_THREADS / 32;

    WindowShape shape = window_shape(rowinfo, lvlptr, winptr, w0);
    window_fetch(stages[0], shape, rowinfo, col, val, lvlptr);
    __pipeline_commit();

    for (int w = w0; w < w1; ++w)
    {
        const WindowStage &st = stages[(w - w0) & 1];
        // this window's copies have landed; everybody is done with the other stage
        __pipeline_wait_prior(0);
        __syncthreads();
        WindowShape ahead = shape;
        if (w + 1 < w1)
        {
            ahead = window_shape(rowinfo, lvlptr, winptr, w + 1);
            window_fetch(stages[(w + 1 - w0) & 1], ahead, rowinfo, col, val, lvlptr);
        }
        __pipeline_commit();

        if (shape.direct)
        {
            for (int r = shape.ra + warp; r < shape.rb; r += nwarps)
                solve_row_all(rowinfo[r], col, val, diag, X, ld, nc, lane);
            __syncthreads();
            shape = ahead;
            continue;
        }

        for (int l = 0; l < shape.lb - shape.la; ++l)
        {
            const int r0 = st.lvl[l] - shape.ra;
            const int rows = st.lvl[l + 1] - shape.ra - r0;
            if (rows >= nwarps / 2)
            {
                // a warp per row, matrix data from the stage
                for (int r = r0 + warp; r < r0 + rows; r += nwarps)
                {
                    const int4 info = st.info[r];
                    for (int c0 = 0; c0 < nc; c0 += 4)
                    {
                        const int ct = min(4, nc - c0);
                        double acc[4] = {0.0, 0.0, 0.0, 0.0};
                        for (int p = info.y - shape.ea + lane; p < info.z - shape.ea; p += 32)
                        {
                            const int j = st.col[p];
                            const double v = st.val[p];
#pragma unroll
                            for (int c = 0; c < 4; ++c)
                                if (c < ct)
                                    acc[c] = fma(v, __ldcg(X + j + (size_t)(c0 + c) * ld), acc[c]);
                        }
#pragma unroll
                        for (int c = 0; c < 4; ++c)
                            if (c < ct)
                            {
#pragma unroll
                                for (int off = 16; off > 0; off >>= 1)
                                    acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], off);
                            }
                        if (lane < ct)
                        {
                            double mine = acc[0];
                            if (lane == 1) mine = acc[1];
                            if (lane == 2) mine = acc[2];
                            if (lane == 3) mine = acc[3];
                            double *x = X + info.x + (size_t)(c0 + lane) * ld;
                            const double r = __ldcg(x) - mine;
                            *x = diag ? r / diag[info.x] : r;
                        }
                    }
                }
                __syncthreads();
                continue;
            }

            // team mode: 16 / pow2(rows) warps per row
            const int shift = 4 - (rows <= 1 ? 0 : 32 - __clz(rows - 1));      // log2(warps per row)
            const int team_warps = 1 << shift;
            const int team = warp >> shift;
            const int part = warp & (team_warps - 1);
            const bool has = team < rows;
            int4 info = make_int4(-1, 0, 0, 0);
            if (has)
                info = st.info[r0 + team];
            const bool leader = has && part == 0;
            const int stride = team_warps * 32;
            for (int c0 = 0; c0 < nc; c0 += 4)
            {
                const int ct = min(4, nc - c0);
                // right-hand side and pivot of this row: do not depend on the sweep
                double rhs = 0.0, d = 1.0;
                if (leader && lane < ct)
                {
                    rhs = __ldcg(X + info.x + (size_t)(c0 + lane) * ld);
                    if (diag)
                        d = diag[info.x];
                }
                double acc[4] = {0.0, 0.0, 0.0, 0.0};
                if (has)
                    for (int p = info.y - shape.ea + part * 32 + lane; p < info.z - shape.ea; p += stride)
                    {
                        const int j = st.col[p];
                        const double v = st.val[p];
#pragma unroll
                        for (int c = 0; c < 4; ++c)
                            if (c < ct)
                                acc[c] = fma(v, __ldcg(X + j + (size_t)(c0 + c) * ld), acc[c]);
                    }
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    if (c < ct)
                    {
#pragma unroll
                        for (int off = 16; off > 0; off >>= 1)
                            acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], off);
                    }
                if (lane == 0)
                {
#pragma unroll
                    for (int c = 0; c < 4; ++c)
                        partial[warp][c] = acc[c];
                }
                __syncthreads();
                if (leader && lane < ct)
                {
                    double sum = 0.0;
                    for (int t = 0; t < team_warps; ++t)
                        sum += partial[warp + t][lane];
                    const double r = rhs - sum;
                    X[info.x + (size_t)(c0 + lane) * ld] = diag ? r / d : r;
                }
                __syncthreads();
            }
        }
        shape = ahead;
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
        seg.win0 = seg.win1 = 0;
        T.segments.push_back(seg);
        l = seg.lvl1;
    }
    // entries per level (strict part), for the windows of the chains
    std::vector<long long> lvl_entries(nlevels, 0);
    for (int i = 0; i < n; ++i)
        lvl_entries[level[i]] += ptr[i + 1] - ptr[i];
    std::vector<int> winptr;
    for (size_t si = 0; si < T.segments.size(); ++si)
    {
        Triangle::Segment &seg = T.segments[si];
        if (!seg.chain)
            continue;
        seg.win0 = (int)winptr.size();
        for (int lv = seg.lvl0; lv < seg.lvl1;)
        {
            const int rows0 = lvlptr[lv + 1] - lvlptr[lv];
            if (lvl_entries[lv] > WIN_ENTRIES || rows0 > WIN_ROWS)
            {
                winptr.push_back(lv | WIN_DIRECT);
                ++lv;
                continue;
            }
            long long entries = 0;
            int rows = 0, e = lv;
            while (e < seg.lvl1 && e - lv < WIN_LEVELS && entries + lvl_entries[e] <= WIN_ENTRIES &&
                   rows + (lvlptr[e + 1] - lvlptr[e]) <= WIN_ROWS)
            {
                entries += lvl_entries[e];
                rows += lvlptr[e + 1] - lvlptr[e];
                ++e;
            }
            winptr.push_back(lv);
            lv = e;
        }
        seg.win1 = (int)winptr.size();
        winptr.push_back(seg.lvl1);          // end marker, shared with nothing
    }

    // entries stored in level order
    std::vector<int4> rowinfo(n);
    std::vector<int> col2(idx.size());
    std::vector<double> val2(val.size());
    int pos = 0;
    for (int r = 0; r < n; ++r)
    {
        const int i = order[r];
        rowinfo[r] = make_int4(i, pos, pos + (ptr[i + 1] - ptr[i]), 0);
        for (int p = ptr[i]; p < ptr[i + 1]; ++p, ++pos)
        {
            col2[pos] = idx[p];
            val2[pos] = val[p];
        }
    }

    int rc;
    if ((rc = upload(rowinfo.data(), (size_t)n * sizeof(int4), (void **)&T.rowinfo))) return rc;
    if ((rc = upload(col2.data(), col2.size() * sizeof(int), (void **)&T.col))) return rc;
    if ((rc = upload(val2.data(), val2.size() * sizeof(double), (void **)&T.val))) return rc;
    if (diag && (rc = upload(diag->data(), (size_t)n * sizeof(double), (void **)&T.diag))) return rc;
    if ((rc = upload(lvlptr.data(), lvlptr.size() * sizeof(int), (void **)&T.lvlptr))) return rc;
    if ((rc = upload(winptr.data(), winptr.size() * sizeof(int), (void **)&T.winptr))) return rc;
    RAILS_CUDA_TRY(cudaStreamSynchronize(ctx().stream));      // host vectors die with the caller
    T.bytes = (size_t)n * 16 + idx.size() * 12 + (diag ? (size_t)n * 8 : 0) + lvlptr.size() * 4;
    return RAILS_OK;
}

void free_triangle(Triangle &T)
{
    cudaFree(T.rowinfo); cudaFree(T.col); cudaFree(T.val); cudaFree(T.diag);
    cudaFree(T.lvlptr); cudaFree(T.winptr);
    T = Triangle();
}

int sweep(const Triangle &T, double *X, long long ld, int nc)
{
    static bool configured = false;
    if (!configured)
    {
        RAILS_CUDA_TRY(cudaFuncSetAttribute(sptrsv_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)(2 * sizeof(WindowStage))));
        configured = true;
    }
    for (size_t s = 0; s < T.segments.size(); ++s)
    {
        const Triangle::Segment &seg = T.segments[s];
        if (seg.chain)
            RAILS_LAUNCH(sptrsv_chain_kernel, 1, CHAIN_THREADS, 2 * sizeof(WindowStage), T.rowinfo, T.col,
                         T.val, T.diag, T.lvlptr, T.winptr, seg.win0, seg.win1, X, ld, nc);
        else
        {
            const int rows = seg.row1 - seg.row0;
            const int warps_per_cta = LEVEL_THREADS / 32;
            int grid = (rows + warps_per_cta - 1) / warps_per_cta;
            grid = std::min(grid, ctx().sm_count * 8);
            RAILS_LAUNCH(sptrsv_level_kernel, grid, LEVEL_THREADS, 0, T.rowinfo, T.col, T.val, T.diag,
                         seg.row0, seg.row1, X, ld, nc);
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
