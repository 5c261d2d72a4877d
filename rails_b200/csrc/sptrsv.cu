// Sparse LU solves on the device: the `A11 \ x` inside every apply of the
// Schur-complement operator (/root/reference/src/SchurOperator.cpp:204-208,
// Amesos_Klu::Solve; matlab/RAILSschur.m:32-33) and the `A \ x` of the inverse
// projection methods (matlab/RAILSschur.m:62-66).
//
// The factors P A Q = L U come from the host factorisation
// (rails_b200/host/SparseLU.cpp) once per operator; here they are kept in HBM
// as four row-oriented triangular matrices (L, U for A^-1; U', L' for A'^-1),
// rows stored in LEVEL ORDER: row i is in level l when the longest dependency
// chain that ends in it has l rows.  A solve is
//     permute in -> sweep the levels of the first factor -> of the second ->
//     permute out,
// in place on an internal work block.
//
// Under a fill-reducing ordering a factor has two very different parts:
//   * the leaves of the elimination tree: a few levels with thousands of short
//     rows.  One launch per level over all SMs, a warp per row, lanes over the
//     row's entries, up to 4 right-hand sides in registers;
//   * the separators near the root: hundreds of consecutive levels with a
//     handful of rows of several hundred entries each, every level waiting for
//     the one before.  Such a run (a "chain") costs one dependent memory round
//     trip per level if it is swept row by row.  It is taken apart instead:
//       1. what the rows of the chain need from OUTSIDE the chain is final before
//          the chain starts: one launch over all SMs subtracts it for all rows
//          of the chain at once;
//       2. what they need from EACH OTHER is swept by ONE CTA, a block barrier
//          between levels, with the chain's part of x held in shared memory
//          (columns renumbered to slots on the host): between two levels stand a
//          shared-memory gather, a warp reduction and two barriers, no global
//          memory.  Descriptors and entries of the chain -- contiguous because of
//          the level order -- are streamed through two shared-memory stages with
//          cp.async, one window of levels ahead.  A level with fewer than 8 rows
//          gives each row a team of 16 / pow2(rows) warps.
// Summation order inside a row is fixed (lane-strided partial sums, xor-shuffle
// tree, warps of a team in order): results are bit-reproducible.
//
// Bound: latency of the level chain for small factors, HBM bandwidth for large
// ones; algorithmic bytes per solve = 12 nnz(L + U) + 16 n c (c right-hand sides).
#include "internal.h"

#include <cuda_pipeline.h>

#include <algorithm>

using rails::ctx;

namespace
{

constexpr int LEVEL_THREADS = 256;         // multi-CTA level kernel: 8 warps
constexpr int CHAIN_THREADS = 512;         // single-CTA chain kernel: 16 warps
constexpr int SMALL_LEVEL_ROWS = 128;      // levels up to this size may join a chain
constexpr int SLOT_ROWS = 2048;            // rows of one chain: its x lives in shared memory
constexpr int CHAIN_COLS = 4;              // right-hand sides per chain launch
// a chain is cut into windows whose matrix data fit one shared-memory stage
constexpr int WIN_ENTRIES = 4096;          // slot (4 B) + value (8 B) per entry: 48 KB
constexpr int WIN_ROWS = 256;              // int4 descriptors: 4 KB
constexpr int WIN_LEVELS = 64;

struct Segment
{
    int lvl0, lvl1;      // levels
    int row0, row1;      // rows in level order
    bool chain;
    int win0, win1;      // windows of a chain
};

// Host-side schedule of one triangle, in the layout the kernels read.
struct TrianglePlan
{
    int n = 0;
    long long nnz = 0;
    int nlevels = 0;
    std::vector<int> lvlptr;        // rows of level l: positions lvlptr[l] .. lvlptr[l+1]
    // main arrays, rows in level order: rowinfo[r] = {row, first entry, end, 0}.
    // A row of a level segment has all its entries here; a row of a chain only
    // those whose column lies outside the chain (col = row id of the column).
    std::vector<int4> rowinfo;
    std::vector<int> col;
    std::vector<double> val;
    // chain arrays: crowinfo[r] = {row, first entry, end, slot}; ccol = slot of
    // the column (position inside the chain)
    std::vector<int4> crowinfo;
    std::vector<int> ccol;
    std::vector<double> cval;
    std::vector<int> winptr;        // first level of window w; each chain's list ends with its last level
    std::vector<int> winshape;      // 8 per entry of winptr: levels la, lb; rows ra, rb; entries ea, eb
    std::vector<Segment> segments;
};

// ptr/idx/val: the STRICT part by rows (entries of row i have col < i when
// `lower`, col > i otherwise).
TrianglePlan plan_triangle(int n, const std::vector<int> &ptr, const std::vector<int> &idx,
                           const std::vector<double> &val, bool lower)
{
    TrianglePlan P;
    P.n = n;
    P.nnz = (long long)idx.size();
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
    P.nlevels = nlevels;
    std::vector<int> &lvlptr = P.lvlptr;
    lvlptr.assign(nlevels + 1, 0);
    std::vector<int> order(n), pos(n);
    for (int i = 0; i < n; ++i)
        ++lvlptr[level[i] + 1];
    for (int l = 0; l < nlevels; ++l)
        lvlptr[l + 1] += lvlptr[l];
    {
        std::vector<int> next(lvlptr.begin(), lvlptr.end() - 1);
        for (int i = 0; i < n; ++i)
        {
            pos[i] = next[level[i]]++;
            order[pos[i]] = i;
        }
    }
    std::vector<long long> lvl_entries(nlevels, 0);
    for (int i = 0; i < n; ++i)
        lvl_entries[level[i]] += ptr[i + 1] - ptr[i];

    // segments: runs of small levels become chains (bounded by the shared-memory
    // slots), every other level is a launch of its own
    auto small = [&](int l) {
        return lvlptr[l + 1] - lvlptr[l] <= SMALL_LEVEL_ROWS && lvl_entries[l] <= WIN_ENTRIES;
    };
    for (int l = 0; l < nlevels;)
    {
        Segment seg;
        seg.lvl0 = l;
        seg.win0 = seg.win1 = 0;
        if (!small(l))
        {
            seg.lvl1 = l + 1;
            seg.chain = false;
        }
        else
        {
            int e = l + 1;
            while (e < nlevels && small(e) && lvlptr[e + 1] - lvlptr[l] <= SLOT_ROWS)
                ++e;
            seg.lvl1 = e;
            // a chain of one level gains nothing over a plain launch
            seg.chain = e - l > 1;
        }
        seg.row0 = lvlptr[seg.lvl0];
        seg.row1 = lvlptr[seg.lvl1];
        P.segments.push_back(seg);
        l = seg.lvl1;
    }

    P.rowinfo.resize(n);
    P.crowinfo.assign(n, make_int4(-1, 0, 0, 0));
    for (size_t si = 0; si < P.segments.size(); ++si)
    {
        Segment &seg = P.segments[si];
        for (int r = seg.row0; r < seg.row1; ++r)
        {
            const int i = order[r];
            const int mbeg = (int)P.col.size(), cbeg = (int)P.ccol.size();
            for (int p = ptr[i]; p < ptr[i + 1]; ++p)
            {
                const int pj = pos[idx[p]];
                if (seg.chain && pj >= seg.row0 && pj < seg.row1)
                {
                    P.ccol.push_back(pj - seg.row0);
                    P.cval.push_back(val[p]);
                }
                else
                {
                    P.col.push_back(idx[p]);
                    P.val.push_back(val[p]);
                }
            }
            P.rowinfo[r] = make_int4(i, mbeg, (int)P.col.size(), 0);
            if (seg.chain)
                P.crowinfo[r] = make_int4(i, cbeg, (int)P.ccol.size(), r - seg.row0);
        }
        if (!seg.chain)
            continue;
        // windows over the chain's own entries
        seg.win0 = (int)P.winptr.size();
        for (int lv = seg.lvl0; lv < seg.lvl1;)
        {
            long long entries = 0;
            int rows = 0, e = lv;
            while (e < seg.lvl1 && e - lv < WIN_LEVELS)
            {
                const int ra = lvlptr[e], rb = lvlptr[e + 1];
                const long long here = P.crowinfo[rb - 1].z - P.crowinfo[ra].y;
                if (e > lv && (entries + here > WIN_ENTRIES || rows + (rb - ra) > WIN_ROWS))
                    break;
                entries += here;
                rows += rb - ra;
                ++e;
            }
            P.winptr.push_back(lv);
            lv = e;
        }
        seg.win1 = (int)P.winptr.size();
        P.winptr.push_back(seg.lvl1);
    }
    P.winshape.assign(P.winptr.size() * 8, 0);
    for (size_t si = 0; si < P.segments.size(); ++si)
        for (int w = P.segments[si].win0; w < P.segments[si].win1; ++w)
        {
            int *sh = &P.winshape[(size_t)w * 8];
            sh[0] = P.winptr[w];
            sh[1] = P.winptr[w + 1];
            sh[2] = lvlptr[sh[0]];
            sh[3] = lvlptr[sh[1]];
            sh[4] = P.crowinfo[sh[2]].y;
            sh[5] = P.crowinfo[sh[3] - 1].z;
        }
    return P;
}

// The sweep as the device runs it, on the host: used to check a plan (slots,
// windows, the split of the chain rows) without a GPU.
void emulate_sweep(const TrianglePlan &P, const double *diag, double *X, long long ld, int nc)
{
    std::vector<double> xs((size_t)SLOT_ROWS);
    for (int c = 0; c < nc; ++c)
    {
        double *x = X + (size_t)c * ld;
        for (size_t si = 0; si < P.segments.size(); ++si)
        {
            const Segment &seg = P.segments[si];
            for (int r = seg.row0; r < seg.row1; ++r)
            {
                const int4 info = P.rowinfo[r];
                double acc = 0.0;
                for (int p = info.y; p < info.z; ++p)
                    acc += P.val[p] * x[P.col[p]];
                const double v = x[info.x] - acc;
                x[info.x] = (!seg.chain && diag) ? v / diag[info.x] : v;
            }
            if (!seg.chain)
                continue;
            for (int w = seg.win0; w < seg.win1; ++w)
                for (int l = P.winptr[w]; l < P.winptr[w + 1]; ++l)
                    for (int r = P.lvlptr[l]; r < P.lvlptr[l + 1]; ++r)
                    {
                        const int4 info = P.crowinfo[r];
                        double acc = 0.0;
                        for (int p = info.y; p < info.z; ++p)
                            acc += P.cval[p] * xs[P.ccol[p]];
                        const double v = x[info.x] - acc;
                        xs[info.w] = x[info.x] = diag ? v / diag[info.x] : v;
                    }
        }
    }
}

struct Triangle
{
    int n = 0;
    long long nnz = 0;
    int nlevels = 0;
    int4 *rowinfo = nullptr;
    int *col = nullptr;
    double *val = nullptr;
    int4 *crowinfo = nullptr;
    int *ccol = nullptr;
    double *cval = nullptr;
    double *diag = nullptr;                // by row; nullptr: unit diagonal
    int *lvlptr = nullptr;
    int *winshape = nullptr;
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

// Rows r0 .. r1 (level order), a warp per row, grid-strided:
//   x_row = (x_row - sum_j T(row, j) x_j) / d_row     (d = 1 when diag is null).
// One level of the leaves, or -- with diag = null -- step 1 of a chain.
__global__ void __launch_bounds__(LEVEL_THREADS)
sptrsv_level_kernel(const int4 *__restrict__ rowinfo, const int *__restrict__ col,
                    const double *__restrict__ val, const double *__restrict__ diag, int r0, int r1,
                    double *X, long long ld, int nc)
{
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * LEVEL_THREADS + threadIdx.x) >> 5;
    const int nwarps = (gridDim.x * LEVEL_THREADS) >> 5;
    for (int r = r0 + warp; r < r1; r += nwarps)
    {
        const int4 info = rowinfo[r];
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
}

// Step 2 of a chain: windows w0 .. w1 walked by one CTA for right-hand sides
// c0 .. c0 + CT (CT <= CHAIN_COLS).
struct WindowStage
{
    int lvl[WIN_LEVELS + 1];     // row offsets of the window's levels
    int4 info[WIN_ROWS];         // {row, first entry, end, slot}
    int col[WIN_ENTRIES];        // slots
    double val[WIN_ENTRIES];
    double rhs[CHAIN_COLS][WIN_ROWS];    // x_row after step 1, final before this kernel starts
    double pivot[WIN_ROWS];
};

struct WindowShape
{
    int la, lb;        // levels
    int ra, rb;        // rows (level order)
    int ea, eb;        // entries
};

struct ChainShared
{
    WindowStage stage[2];
    double x[CHAIN_COLS][SLOT_ROWS];
};

// the shape of window w, precomputed on the host: two 16-byte loads, no chain
__device__ __forceinline__ WindowShape window_shape(const int *__restrict__ winshape, int w)
{
    const int4 a = *reinterpret_cast<const int4 *>(winshape + 8 * (size_t)w);
    const int4 b = *reinterpret_cast<const int4 *>(winshape + 8 * (size_t)w + 4);
    WindowShape s;
    s.la = a.x; s.lb = a.y; s.ra = a.z; s.rb = a.w;
    s.ea = b.x; s.eb = b.y;
    return s;
}

// Everything a window needs that does not depend on the sweep: level offsets, row
// descriptors, entries, and -- final since step 1 -- the right-hand sides and pivots
// of its rows.
__device__ __forceinline__ void window_fetch(WindowStage &st, const WindowShape &s,
                                             const int4 *__restrict__ crowinfo,
                                             const int *__restrict__ ccol,
                                             const double *__restrict__ cval,
                                             const int *__restrict__ lvlptr,
                                             const double *__restrict__ diag, const double *X,
                                             long long ld, int c0, int ct)
{
    for (int i = threadIdx.x; i < s.rb - s.ra; i += CHAIN_THREADS)
    {
        const int4 info = crowinfo[s.ra + i];
        st.info[i] = info;
        for (int c = 0; c < ct; ++c)
            __pipeline_memcpy_async(&st.rhs[c][i], X + info.x + (size_t)(c0 + c) * ld, 8);
        if (diag)
            __pipeline_memcpy_async(&st.pivot[i], diag + info.x, 8);
    }
    for (int i = threadIdx.x; i <= s.lb - s.la; i += CHAIN_THREADS)
        __pipeline_memcpy_async(&st.lvl[i], lvlptr + s.la + i, 4);
    for (int i = threadIdx.x; i < s.eb - s.ea; i += CHAIN_THREADS)
    {
        __pipeline_memcpy_async(&st.col[i], ccol + s.ea + i, 4);
        __pipeline_memcpy_async(&st.val[i], cval + s.ea + i, 8);
    }
}

// CT = right-hand sides of this launch (1 .. CHAIN_COLS): a template parameter, so
// that one right-hand side does not pay the issue slots of four
template<int CT>
__global__ void __launch_bounds__(CHAIN_THREADS, 1)
sptrsv_chain_kernel(const int4 *__restrict__ crowinfo, const int *__restrict__ ccol,
                    const double *__restrict__ cval, const double *__restrict__ diag,
                    const int *__restrict__ lvlptr, const int *__restrict__ winshape, int w0, int w1,
                    double *X, long long ld, int c0)
{
    constexpr int ct = CT;
    extern __shared__ __align__(16) unsigned char chain_smem[];
    ChainShared &sh = *reinterpret_cast<ChainShared *>(chain_smem);
    __shared__ double partial[CHAIN_THREADS / 32][CT];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    constexpr int nwarps = CHAIN_THREADS / 32;

    WindowShape shape = window_shape(winshape, w0);
    window_fetch(sh.stage[0], shape, crowinfo, ccol, cval, lvlptr, diag, X, ld, c0, ct);
    __pipeline_commit();
    WindowShape ahead = window_shape(winshape, min(w0 + 1, w1 - 1));

    for (int w = w0; w < w1; ++w)
    {
        const WindowStage &st = sh.stage[(w - w0) & 1];
        // this window's copies have landed; everybody is done with the other stage
        __pipeline_wait_prior(0);
        __syncthreads();
        // shapes are loaded two windows ahead, the data one window ahead
        const WindowShape ahead2 = window_shape(winshape, min(w + 2, w1 - 1));
        if (w + 1 < w1)
            window_fetch(sh.stage[(w + 1 - w0) & 1], ahead, crowinfo, ccol, cval, lvlptr, diag, X, ld, c0, ct);
        __pipeline_commit();

        for (int l = 0; l < shape.lb - shape.la; ++l)
        {
            const int r0 = st.lvl[l] - shape.ra;
            const int rows = st.lvl[l + 1] - shape.ra - r0;
            if (rows >= nwarps / 2)
            {
                // a warp per row
                for (int r = r0 + warp; r < r0 + rows; r += nwarps)
                {
                    const int4 info = st.info[r];
                    double rhs = 0.0, d = 1.0;
                    if (lane < ct)
                    {
                        rhs = st.rhs[lane][r];
                        if (diag)
                            d = st.pivot[r];
                    }
                    double acc[CT];
#pragma unroll
                    for (int c = 0; c < CT; ++c)
                        acc[c] = 0.0;
                    for (int p = info.y - shape.ea + lane; p < info.z - shape.ea; p += 32)
                    {
                        const int slot = st.col[p];
                        const double v = st.val[p];
#pragma unroll
                        for (int c = 0; c < CT; ++c)
                            acc[c] = fma(v, sh.x[c][slot], acc[c]);
                    }
#pragma unroll
                    for (int off = 16; off > 0; off >>= 1)
#pragma unroll
                        for (int c = 0; c < CT; ++c)
                            acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], off);
                    if (lane < ct)
                    {
                        double mine = acc[0];
#pragma unroll
                        for (int c = 1; c < CT; ++c)
                            if (lane == c)
                                mine = acc[c];
                        const double r1 = rhs - mine;
                        const double xv = diag ? r1 / d : r1;
                        X[info.x + (size_t)(c0 + lane) * ld] = xv;
                        sh.x[lane][info.w] = xv;
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
            // right-hand side and pivot of this row: do not depend on the sweep
            double rhs = 0.0, d = 1.0;
            if (leader && lane < ct)
            {
                rhs = st.rhs[lane][r0 + team];
                if (diag)
                    d = st.pivot[r0 + team];
            }
            double acc[CT];
#pragma unroll
            for (int c = 0; c < CT; ++c)
                acc[c] = 0.0;
            if (has)
                for (int p = info.y - shape.ea + part * 32 + lane; p < info.z - shape.ea; p += team_warps * 32)
                {
                    const int slot = st.col[p];
                    const double v = st.val[p];
#pragma unroll
                    for (int c = 0; c < CT; ++c)
                        acc[c] = fma(v, sh.x[c][slot], acc[c]);
                }
#pragma unroll
            for (int off = 16; off > 0; off >>= 1)
#pragma unroll
                for (int c = 0; c < CT; ++c)
                    acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], off);
            if (lane == 0)
            {
#pragma unroll
                for (int c = 0; c < CT; ++c)
                    partial[warp][c] = acc[c];
            }
            __syncthreads();
            if (leader && lane < ct)
            {
                double sum = 0.0;
                for (int t = 0; t < team_warps; ++t)
                    sum += partial[warp + t][lane];
                const double r1 = rhs - sum;
                const double xv = diag ? r1 / d : r1;
                X[info.x + (size_t)(c0 + lane) * ld] = xv;
                sh.x[lane][info.w] = xv;
            }
            __syncthreads();
        }
        shape = ahead;
        ahead = ahead2;
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

int build_triangle(int n, const std::vector<int> &ptr, const std::vector<int> &idx,
                   const std::vector<double> &val, const std::vector<double> *diag, bool lower,
                   Triangle &T)
{
    const TrianglePlan P = plan_triangle(n, ptr, idx, val, lower);
    T.n = n;
    T.nnz = P.nnz;
    T.nlevels = P.nlevels;
    T.segments = P.segments;
    int rc;
    if ((rc = upload(P.rowinfo.data(), P.rowinfo.size() * sizeof(int4), (void **)&T.rowinfo))) return rc;
    if ((rc = upload(P.col.data(), P.col.size() * sizeof(int), (void **)&T.col))) return rc;
    if ((rc = upload(P.val.data(), P.val.size() * sizeof(double), (void **)&T.val))) return rc;
    if ((rc = upload(P.crowinfo.data(), P.crowinfo.size() * sizeof(int4), (void **)&T.crowinfo))) return rc;
    if ((rc = upload(P.ccol.data(), P.ccol.size() * sizeof(int), (void **)&T.ccol))) return rc;
    if ((rc = upload(P.cval.data(), P.cval.size() * sizeof(double), (void **)&T.cval))) return rc;
    if (diag && (rc = upload(diag->data(), (size_t)n * sizeof(double), (void **)&T.diag))) return rc;
    if ((rc = upload(P.lvlptr.data(), P.lvlptr.size() * sizeof(int), (void **)&T.lvlptr))) return rc;
    if ((rc = upload(P.winshape.data(), P.winshape.size() * sizeof(int), (void **)&T.winshape))) return rc;
    RAILS_CUDA_TRY(cudaStreamSynchronize(ctx().stream));      // the plan dies with this call
    T.bytes = (size_t)n * 32 + idx.size() * 12 + (diag ? (size_t)n * 8 : 0) +
              (P.lvlptr.size() + P.winshape.size()) * 4;
    return RAILS_OK;
}

void free_triangle(Triangle &T)
{
    cudaFree(T.rowinfo); cudaFree(T.col); cudaFree(T.val);
    cudaFree(T.crowinfo); cudaFree(T.ccol); cudaFree(T.cval);
    cudaFree(T.diag); cudaFree(T.lvlptr); cudaFree(T.winshape);
    T = Triangle();
}

int launches_of(const Triangle &T, int nc)
{
    int launches = 0;
    for (size_t s = 0; s < T.segments.size(); ++s)
        launches += T.segments[s].chain ? 1 + (nc + CHAIN_COLS - 1) / CHAIN_COLS : 1;
    return launches;
}

int sweep(const Triangle &T, double *X, long long ld, int nc)
{
    static int configured_device = -1;          // the attribute is per device
    if (configured_device != ctx().device)
    {
        RAILS_CUDA_TRY(cudaFuncSetAttribute(sptrsv_chain_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)sizeof(ChainShared)));
        RAILS_CUDA_TRY(cudaFuncSetAttribute(sptrsv_chain_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)sizeof(ChainShared)));
        RAILS_CUDA_TRY(cudaFuncSetAttribute(sptrsv_chain_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)sizeof(ChainShared)));
        RAILS_CUDA_TRY(cudaFuncSetAttribute(sptrsv_chain_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)sizeof(ChainShared)));
        configured_device = ctx().device;
    }
    const int warps_per_cta = LEVEL_THREADS / 32;
    for (size_t s = 0; s < T.segments.size(); ++s)
    {
        const Segment &seg = T.segments[s];
        const int rows = seg.row1 - seg.row0;
        int grid = (rows + warps_per_cta - 1) / warps_per_cta;
        grid = std::max(1, std::min(grid, ctx().sm_count * 8));
        // a level of the leaves, or what the chain's rows need from outside the chain
        RAILS_LAUNCH(sptrsv_level_kernel, grid, LEVEL_THREADS, 0, T.rowinfo, T.col, T.val,
                     seg.chain ? (const double *)nullptr : T.diag, seg.row0, seg.row1, X, ld, nc);
        if (!seg.chain)
            continue;
        for (int c0 = 0; c0 < nc; c0 += CHAIN_COLS)
        {
            switch (std::min(CHAIN_COLS, nc - c0))
            {
            case 1:
                RAILS_LAUNCH(sptrsv_chain_kernel<1>, 1, CHAIN_THREADS, sizeof(ChainShared), T.crowinfo, T.ccol,
                             T.cval, T.diag, T.lvlptr, T.winshape, seg.win0, seg.win1, X, ld, c0);
                break;
            case 2:
                RAILS_LAUNCH(sptrsv_chain_kernel<2>, 1, CHAIN_THREADS, sizeof(ChainShared), T.crowinfo, T.ccol,
                             T.cval, T.diag, T.lvlptr, T.winshape, seg.win0, seg.win1, X, ld, c0);
                break;
            case 3:
                RAILS_LAUNCH(sptrsv_chain_kernel<3>, 1, CHAIN_THREADS, sizeof(ChainShared), T.crowinfo, T.ccol,
                             T.cval, T.diag, T.lvlptr, T.winshape, seg.win0, seg.win1, X, ld, c0);
                break;
            default:
                RAILS_LAUNCH(sptrsv_chain_kernel<4>, 1, CHAIN_THREADS, sizeof(ChainShared), T.crowinfo, T.ccol,
                             T.cval, T.diag, T.lvlptr, T.winshape, seg.win0, seg.win1, X, ld, c0);
                break;
            }
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

// Host-only: the four schedules built as rails_lu_create builds them, and one
// solve executed the way the kernels traverse them (see emulate_sweep).
extern "C" int rails_lu_plan_emulate(int n, const int *Lp, const int *Li, const double *Lx,
                                     const int *Up, const int *Ui, const double *Ux,
                                     const int *pinv, const int *q, int trans, double *x, int nrhs,
                                     long long ld, long long *stats)
{
    if (n < 0 || !Lp || !Up || !pinv || !q || !x || nrhs < 0 || ld < n)
        return rails::fail(RAILS_ERR_ARG, "rails_lu_plan_emulate: bad argument");
    std::vector<int> ptr, idx;
    std::vector<double> val, diag(n, 0.0);
    TrianglePlan first, second;
    if (!trans)
    {
        transpose_strict(n, Lp, Li, Lx, ptr, idx, val, nullptr);
        first = plan_triangle(n, ptr, idx, val, true);
        transpose_strict(n, Up, Ui, Ux, ptr, idx, val, &diag);
        second = plan_triangle(n, ptr, idx, val, false);
    }
    else
    {
        transpose_strict(n, Up, Ui, Ux, ptr, idx, val, &diag);      // for the diagonal
        copy_strict(n, Up, Ui, Ux, ptr, idx, val);
        first = plan_triangle(n, ptr, idx, val, true);
        copy_strict(n, Lp, Li, Lx, ptr, idx, val);
        second = plan_triangle(n, ptr, idx, val, false);
    }
    std::vector<double> w((size_t)n * (size_t)std::max(nrhs, 1));
    for (int c = 0; c < nrhs; ++c)
        for (int i = 0; i < n; ++i)
        {
            if (!trans)
                w[(size_t)c * n + pinv[i]] = x[(size_t)c * ld + i];
            else
                w[(size_t)c * n + i] = x[(size_t)c * ld + q[i]];
        }
    // non-transposed: L is unit, U carries the diagonal; transposed: U' first (diagonal), then L'
    emulate_sweep(first, trans ? diag.data() : nullptr, w.data(), n, nrhs);
    emulate_sweep(second, trans ? nullptr : diag.data(), w.data(), n, nrhs);
    for (int c = 0; c < nrhs; ++c)
        for (int i = 0; i < n; ++i)
        {
            if (!trans)
                x[(size_t)c * ld + q[i]] = w[(size_t)c * n + i];
            else
                x[(size_t)c * ld + i] = w[(size_t)c * n + pinv[i]];
        }
    if (stats)
    {
        const TrianglePlan *plans[2] = {&first, &second};
        for (int t = 0; t < 2; ++t)
        {
            long long chains = 0, windows = 0, chain_rows = 0;
            for (size_t s2 = 0; s2 < plans[t]->segments.size(); ++s2)
                if (plans[t]->segments[s2].chain)
                {
                    ++chains;
                    windows += plans[t]->segments[s2].win1 - plans[t]->segments[s2].win0;
                    chain_rows += plans[t]->segments[s2].row1 - plans[t]->segments[s2].row0;
                }
            stats[6 * t + 0] = plans[t]->nlevels;
            stats[6 * t + 1] = (long long)plans[t]->segments.size();
            stats[6 * t + 2] = chains;
            stats[6 * t + 3] = windows;
            stats[6 * t + 4] = chain_rows;
            stats[6 * t + 5] = (long long)plans[t]->ccol.size();
        }
    }
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
    return 2 + (trans ? launches_of(F->Ut, 1) + launches_of(F->Lt, 1)
                      : launches_of(F->L, 1) + launches_of(F->U, 1));
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
