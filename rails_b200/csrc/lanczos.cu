// Fused kernels for one step of the residual Lanczos recurrence
// (Solver::resid_lanczos, /root/reference/src/LyapunovSolver.hpp:384-434):
//
//     g    = [V AV B]' q                         rails_mv_multi_dot
//     next = V z_V + AV z_AV + B y ; a = next'q  rails_mv_multi_gemv
//     next -= a q + b q_prev ; b' = |next|       rails_mv_lanczos_update
//
// The reference issues these as ~19 wrapper calls per step (three dots, three
// V*C products with temporaries, two scaled copies, two axpys, a norm, a
// scale), i.e. m (32k + 16n + 264) bytes; fused, a step moves
// m (16 (2k + n) + 64) bytes in three launches: every column of V, AV and B
// is read exactly twice (once per kernel: the coefficients of the second pass
// depend on the complete result of the first), q three times, `next` is
// written twice.  All three are pure HBM streams (<= 0.25 flop/B).
//
// Reductions are deterministic: per-warp shuffle trees, per-CTA partials in
// global memory, a second kernel adds the partials in a fixed order; with a
// communicator the small result is all-reduced over the ranks (NCCL).
//
// Round 2.  The device-resident chain (rails_mv_lanczos_run) runs a step as ONE kernel,
// lanczos_step_kernel: a row tile of [V AV B] is staged once in shared memory by TMA
// and serves both the new vector (row-local) and the next projection, so every column
// is read ONCE per step, m (8 (2k + n) + 32) bytes; alpha comes out of the finish
// kernel as g . coef, so there is one cross-GPU reduction per step.  The three entry
// points above remain as the "step" mode and for shapes the step kernel declines
// (more than 96 block columns).  DESIGN.md section 4, profiles/r02_lanczos_step.md.
#include "internal.h"
#include "tma.cuh"

#include <cuda.h>

#include <cmath>
#include <cstring>

using rails::ctx;

namespace
{

constexpr int MAX_SEGS = 4;
constexpr int MD_THREADS = 256;
constexpr int MD_WARPS = MD_THREADS / 32;
constexpr int MD_RP = 4;                         // double2 per thread and column
constexpr long long MD_TILE = 512LL * MD_RP;     // rows per CTA iteration
constexpr int MD_U = 4;                          // columns in flight
constexpr int MD_MAXCOLS = 2040;                 // 128 KB of per-warp accumulators; + 1 fits one mailbox exchange (2048)

int lz_attr();      // defined with the chain's scratch management below

struct Segs
{
    const double *p[MAX_SEGS];
    long long ld[MAX_SEGS];
    int n[MAX_SEGS];
    int nseg;
};

template<bool FULL>
__device__ __forceinline__ double2 ld2(const double *__restrict__ p, long long r, long long rows)
{
    if (FULL)
        return __ldg(reinterpret_cast<const double2 *>(p + r));
    double2 v = make_double2(0.0, 0.0);
    if (r < rows) v.x = __ldg(p + r);
    if (r + 1 < rows) v.y = __ldg(p + r + 1);
    return v;
}

// Fused update of the device-resident recurrence (rails_mv_lanczos_run): the
// column being projected is first formed as
//     w_i = raw - alpha (s w_{i-1}) - beta q_{i-2},   s = 1 / |w_{i-1}|, beta = |w_{i-1}|
// (src/LyapunovSolver.hpp:411-413 and the `next /= beta` of :431 for the
// PREVIOUS column, which is normalised in place in the same pass), and its
// squared norm is accumulated as one more output.
struct LanczosUpdate
{
    double *prev;            // w_{i-1}, un-normalised on entry, normalised on exit; NULL: no update
    const double *prev2;     // q_{i-2} (normalised) or NULL
    double alpha, s, beta;
};

template<bool FULL, bool UPD>
__device__ __forceinline__ void md_tile(const Segs &S, double *__restrict__ q, double qscale,
                                        bool write_q, long long rows, long long base,
                                        double *wacc, int lane, const LanczosUpdate &up, int total)
{
    // q tile: scaled on the fly (the pending `q /= beta` of the previous
    // step, :431) and written back once
    double2 qv[MD_RP];
    if (UPD)
    {
        double nrm = 0.0;
#pragma unroll
        for (int v = 0; v < MD_RP; ++v)
        {
            const long long r = base + v * 512;
            double2 x = make_double2(0.0, 0.0), w = make_double2(0.0, 0.0), p = make_double2(0.0, 0.0);
            if (FULL)
            {
                x = *reinterpret_cast<const double2 *>(q + r);
                w = *reinterpret_cast<const double2 *>(up.prev + r);
                if (up.prev2) p = __ldg(reinterpret_cast<const double2 *>(up.prev2 + r));
            }
            else
            {
                if (r < rows) { x.x = q[r]; w.x = up.prev[r]; if (up.prev2) p.x = up.prev2[r]; }
                if (r + 1 < rows) { x.y = q[r + 1]; w.y = up.prev[r + 1]; if (up.prev2) p.y = up.prev2[r + 1]; }
            }
            w.x = __dmul_rn(w.x, up.s);
            w.y = __dmul_rn(w.y, up.s);
            x.x = __dadd_rn(x.x, -__dmul_rn(up.alpha, w.x));
            x.y = __dadd_rn(x.y, -__dmul_rn(up.alpha, w.y));
            if (up.prev2)
            {
                x.x = __dadd_rn(x.x, -__dmul_rn(up.beta, p.x));
                x.y = __dadd_rn(x.y, -__dmul_rn(up.beta, p.y));
            }
            if (FULL)
            {
                *reinterpret_cast<double2 *>(q + r) = x;
                *reinterpret_cast<double2 *>(up.prev + r) = w;
            }
            else
            {
                if (r < rows) { q[r] = x.x; up.prev[r] = w.x; }
                if (r + 1 < rows) { q[r + 1] = x.y; up.prev[r + 1] = w.y; }
            }
            qv[v] = x;
            nrm = fma(x.x, x.x, nrm);
            nrm = fma(x.y, x.y, nrm);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
            nrm += __shfl_xor_sync(0xffffffffu, nrm, o);
        if (lane == 0) wacc[total] += nrm;
    }
    else
#pragma unroll
    for (int v = 0; v < MD_RP; ++v)
    {
        const long long r = base + v * 512;
        if (FULL)
            qv[v] = *reinterpret_cast<const double2 *>(q + r);
        else
        {
            qv[v] = make_double2(0.0, 0.0);
            if (r < rows) qv[v].x = q[r];
            if (r + 1 < rows) qv[v].y = q[r + 1];
        }
        if (write_q)
        {
            qv[v].x *= qscale;
            qv[v].y *= qscale;
            if (FULL)
                *reinterpret_cast<double2 *>(q + r) = qv[v];
            else
            {
                if (r < rows) q[r] = qv[v].x;
                if (r + 1 < rows) q[r + 1] = qv[v].y;
            }
        }
    }

    int off = 0;
#pragma unroll
    for (int s = 0; s < MAX_SEGS; ++s)
    {
        if (s >= S.nseg) break;
        const double *__restrict__ X = S.p[s];
        const long long ldx = S.ld[s];
        const int ac = S.n[s];
        for (int j = 0; j < ac; j += MD_U)
        {
            double2 xv[MD_U][MD_RP];
#pragma unroll
            for (int u = 0; u < MD_U; ++u)
            {
                const int jj = (j + u < ac) ? (j + u) : (ac - 1);
#pragma unroll
                for (int v = 0; v < MD_RP; ++v)
                    xv[u][v] = ld2<FULL>(X + (size_t)jj * ldx, base + v * 512, rows);
            }
            double p[MD_U];
#pragma unroll
            for (int u = 0; u < MD_U; ++u)
            {
                double acc = 0.0;
#pragma unroll
                for (int v = 0; v < MD_RP; ++v)
                {
                    acc = fma(xv[u][v].x, qv[v].x, acc);
                    acc = fma(xv[u][v].y, qv[v].y, acc);
                }
                p[u] = acc;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1)
#pragma unroll
                for (int u = 0; u < MD_U; ++u)
                    p[u] += __shfl_xor_sync(0xffffffffu, p[u], o);
            if (lane == 0)
            {
#pragma unroll
                for (int u = 0; u < MD_U; ++u)
                    if (j + u < ac)
                        wacc[off + j + u] += p[u];
            }
        }
        off += ac;
    }
}

// partial[blockIdx.x][c] = sum over this CTA's row tiles of X_c . (qscale q)
// qscale_sq != NULL: the scale is 1 / sqrt(|*qscale_sq|), read from device memory
// (the squared norm the previous step left there), and q is written back.
__global__ void __launch_bounds__(MD_THREADS)
multi_dot_kernel(Segs S, int total, double *__restrict__ q, double qscale, int write_q,
                 long long rows, double *__restrict__ partial, const double *__restrict__ qscale_sq)
{
    extern __shared__ double sacc[];
    if (qscale_sq)
    {
        qscale = 1.0 / sqrt(fabs(*qscale_sq));
        write_q = 1;
    }
    for (int i = threadIdx.x; i < total * MD_WARPS; i += MD_THREADS)
        sacc[i] = 0.0;
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    double *wacc = sacc + warp * total;
    const long long ntiles = (rows + MD_TILE - 1) / MD_TILE;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x)
    {
        const long long base = tile * MD_TILE + 2 * threadIdx.x;
        const LanczosUpdate none = {nullptr, nullptr, 0.0, 0.0, 0.0};
        if ((tile + 1) * MD_TILE <= rows)
            md_tile<true, false>(S, q, qscale, write_q != 0, rows, base, wacc, lane, none, total);
        else
            md_tile<false, false>(S, q, qscale, write_q != 0, rows, base, wacc, lane, none, total);
    }
    __syncthreads();
    double *out = partial + (size_t)blockIdx.x * total;
    for (int i = threadIdx.x; i < total; i += MD_THREADS)
    {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < MD_WARPS; ++w)
            s += sacc[w * total + i];
        out[i] = s;
    }
}

// Step kernel A of the device-resident recurrence:
//   partial[blockIdx.x][0 .. total)  = [X_0 X_1 ..]' w      (w NOT normalised)
//   partial[blockIdx.x][total]       = |w|^2
// where w = col (i = 0) or the fused update of LanczosUpdate with alpha and
// |w_{i-1}|^2 read from device memory.
__global__ void __launch_bounds__(MD_THREADS)
lanczos_dot_kernel(Segs S, int total, double *__restrict__ col, double *prev, const double *prev2,
                   const double *__restrict__ alpha_dev, const double *__restrict__ prev_sq_dev,
                   long long rows, double *__restrict__ partial)
{
    extern __shared__ double sacc[];
    const int nout = total + 1;
    for (int i = threadIdx.x; i < nout * MD_WARPS; i += MD_THREADS)
        sacc[i] = 0.0;
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    double *wacc = sacc + warp * nout;
    LanczosUpdate up = {prev, prev2, 0.0, 0.0, 0.0};
    if (prev)
    {
        const double nrm = sqrt(fabs(*prev_sq_dev));
        up.alpha = *alpha_dev;
        up.s = 1.0 / nrm;
        up.beta = nrm;
    }
    const long long ntiles = (rows + MD_TILE - 1) / MD_TILE;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x)
    {
        const long long base = tile * MD_TILE + 2 * threadIdx.x;
        const bool full = (tile + 1) * MD_TILE <= rows;
        if (prev)
        {
            if (full) md_tile<true, true>(S, col, 1.0, false, rows, base, wacc, lane, up, total);
            else      md_tile<false, true>(S, col, 1.0, false, rows, base, wacc, lane, up, total);
        }
        else
        {
            // first column: plain projection; its squared norm from the q tile
            if (full) md_tile<true, false>(S, col, 1.0, false, rows, base, wacc, lane, up, total);
            else      md_tile<false, false>(S, col, 1.0, false, rows, base, wacc, lane, up, total);
            double nrm = 0.0;
#pragma unroll
            for (int v = 0; v < MD_RP; ++v)
            {
                const long long r = base + v * 512;
                if (r < rows) nrm = fma(col[r], col[r], nrm);
                if (r + 1 < rows) nrm = fma(col[r + 1], col[r + 1], nrm);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1)
                nrm += __shfl_xor_sync(0xffffffffu, nrm, o);
            if (lane == 0) wacc[total] += nrm;
        }
    }
    __syncthreads();
    double *out = partial + (size_t)blockIdx.x * nout;
    for (int i = threadIdx.x; i < nout; i += MD_THREADS)
    {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < MD_WARPS; ++w)
            s += sacc[w * nout + i];
        out[i] = s;
    }
}

// ---- one kernel per Lanczos step -----------------------------------------------------
// With alpha_i available BEFORE the vector of step i+1 is formed (alpha = g . coef in
// the finish of step i, exchange.cu), everything between two reductions is row-local:
//     raw_{i+1}[r] = sum_b X_b[r, :] coef_b                       (kernel B of the chain)
//     q_i[r]       = w_i[r] / |w_i|                                (written back in place)
//     w_{i+1}[r]   = raw_{i+1}[r] - rnd(alpha_i q_i[r]) - rnd(|w_i| q_{i-1}[r])   (:411-413)
//     partial g   += X_b[r, :]' w_{i+1}[r],  |w_{i+1}|^2           (kernel A of step i+1)
// so a row tile of [V|MV, AV, B] is brought from HBM ONCE per step instead of twice:
// 8 m (2k + n) + 32 m bytes per step against 16 m (2k + n) + 56 m.  A CTA per SM walks
// the tiles of TR rows; one thread moves the NEXT tile into the other shared-memory
// stage with one 1-D bulk copy per column (cp.async.bulk + mbarrier: the TMA engine,
// no registers involved) while the CTA works on the current one: phase 1, a thread per
// row, the fma chain over all columns in block order (the arithmetic of
// multi_gemv_kernel) and the update (the arithmetic of md_tile<., true>); phase 2, a
// warp per column, the inner product with the new vector out of the same staged tile.
// Column j is always handled by warp j mod 8 and accumulated in its own slot: no
// atomics, a summation order fixed by (TR, grid) alone.
constexpr int LS_MAX_WARPS = 16;                   // consumer warps (32 rows of a tile each)
constexpr int LS_MAX_STAGES = 6;
constexpr size_t LS_SMEM_BUDGET = 216 * 1024;      // of the 227 KB a CTA may have
constexpr int LS_ACC_COLS = 24;                    // accumulator columns of a warp (96 registers with 17 warps)
constexpr int LS_WIDE_COLS = 50;                   // accumulator columns of a warp with 9 warps (168 registers)
constexpr int LS_MAX_GROUPS = 4;                   // column groups (warps sharing 32 rows)
constexpr int LS_MAX_COLS = LS_MAX_GROUPS * LS_ACC_COLS;      // block columns (2k + n) covered
constexpr int LS_BOX_ROWS = 128;                   // rows of a TMA box (1 KB per column)
constexpr int LS_BOX_COLS = 8;                     // columns of a TMA box: 8 KB per request

// The tile in shared memory: for each block, ceil(n_b / bc_b) boxes of [bc_b][128]
// doubles (bc_b = min(n_b, 8) columns: a block is padded to whole boxes), then one box
// of two columns [q_{i-1}, w_i] of the Lanczos basis; a tile of TR rows is TR / 128 such
// slabs.
struct StepLayout
{
    int nseg;
    int n[MAX_SEGS];        // columns of block b
    int bc[MAX_SEGS];       // box columns of block b
    int off[MAX_SEGS];      // first shared-memory column of block b
    int total;              // sum of n
    int padded;             // shared-memory columns of all blocks
};
struct StepMaps
{
    CUtensorMap m[MAX_SEGS + 1];    // the blocks, then the Lanczos basis Q
};

inline StepLayout ls_layout(const int *n, int nseg)
{
    StepLayout L;
    L.nseg = nseg;
    L.total = L.padded = 0;
    for (int b = 0; b < MAX_SEGS; ++b)
    {
        L.n[b] = b < nseg ? n[b] : 0;
        // box columns: 8 (padding costs shared memory only: columns past the block's
        // end are zero-filled without traffic); narrower boxes that pad less only for
        // blocks too wide to fit otherwise (more, smaller requests: 0.76 -> 0.89 ms at
        // 15 columns when tried everywhere)
        L.bc[b] = std::max(1, std::min(L.n[b], LS_BOX_COLS));
        for (int bc = LS_BOX_COLS; bc >= 4 && L.n[b] > 40; --bc)
            if ((L.n[b] + bc - 1) / bc * bc < (L.n[b] + L.bc[b] - 1) / L.bc[b] * L.bc[b])
                L.bc[b] = bc;
        L.off[b] = L.padded;
        L.total += L.n[b];
        L.padded += (L.n[b] + L.bc[b] - 1) / L.bc[b] * L.bc[b];
    }
    return L;
}
constexpr int LS_WSUM_DOUBLES = 8 * (LS_WIDE_COLS + 1);       // >= 16 * (LS_ACC_COLS + 1): per-warp sums

inline int ls_groups(int total);

inline size_t ls_smem_bytes(StepLayout const &L, int tr, int stages)
{
    // stages [padded + 2][tr] | coef[total, even] | per-warp sums | partial chain sums
    // [2][groups][tr] (several column groups only) | tables (ints)
    const int groups = ls_groups(L.total);
    return ((size_t)stages * (L.padded + 2) * tr + (size_t)((L.total + 1) & ~1) + (size_t)LS_WSUM_DOUBLES +
            (groups > 1 ? (size_t)2 * groups * tr : 0)) * sizeof(double) +
           (size_t)(L.total + 64 * 3 + 2) * sizeof(int) + 128;
}

// Column groups of a row group's warps: one warp covers up to 50 columns with 8
// consumer warps per CTA (168 registers), beyond that 24 columns per warp with 16.
inline int ls_groups(int total)
{
    static const int forced = getenv("RAILS_LS_GROUPS") ? atoi(getenv("RAILS_LS_GROUPS")) : 0;
    if (forced > 0 && forced <= LS_MAX_GROUPS && forced * LS_ACC_COLS >= total)
        return forced;
    return total <= LS_WIDE_COLS ? 1 : (total + LS_ACC_COLS - 1) / LS_ACC_COLS;
}
// Tile height and pipeline depth: candidates in order of preference (measured,
// profiles/r02_lanczos_step.md); the first that fits shared memory wins.  tr = 0: not
// applicable (more than LS_MAX_COLS columns, or no room).
inline void ls_shape(StepLayout const &L, int *tr_out, int *stages_out)
{
    *tr_out = 0;
    *stages_out = 0;
    if (L.total > LS_MAX_COLS || L.total < 1)
        return;
    static int forced_tr = -1, forced_st = -1;   // RAILS_LS_TILE_ROWS / RAILS_LS_STAGES: experiments
    if (forced_tr < 0)
    {
        const char *e = getenv("RAILS_LS_TILE_ROWS");
        forced_tr = e ? atoi(e) : 0;
        e = getenv("RAILS_LS_STAGES");
        forced_st = e ? atoi(e) : 0;
    }
    static const int cand[][2] = {{512, 3}, {512, 2}, {384, 3}, {384, 2}, {256, 3}, {256, 2}, {128, 4}, {128, 3}, {128, 2}};
    for (auto const &cd : cand)
    {
        const int tr = forced_tr > 0 ? forced_tr : cd[0], st = forced_st > 0 ? forced_st : cd[1];
        // column groups x row groups <= 16 consumer warps (8 for the wide instances)
        const int groups = ls_groups(L.total);
        const int max_rows = groups == 1 ? (L.total <= LS_ACC_COLS ? 512 : 256) : 32 * (LS_MAX_WARPS / groups);
        if (tr % LS_BOX_ROWS == 0 && tr <= max_rows && st >= 2 && st <= LS_MAX_STAGES &&
            ls_smem_bytes(L, tr, st) <= LS_SMEM_BUDGET)
        {
            *tr_out = tr;
            *stages_out = st;
            return;
        }
        if (forced_tr > 0 && forced_st > 0)
            return;
    }
}

// Consumer warp w = cg * nrg + rg: row group rg (rows rg * 32 + lane of every tile) and
// column group cg (block columns [cg * CACC, (cg + 1) * CACC)).  With one column group a
// warp does everything for its rows.  With several, each warp of a row group sums the
// chain over ITS columns, the partial sums meet in shared memory behind a named barrier
// of just those warps, every warp forms the new entries, the warp with cg = 0 writes
// them, and each accumulates the products of its columns: shared memory is read twice
// whatever the split.  The last warp is the producer.  Registers: CACC accumulators per
// lane, 96 for 17 warps (CACC <= 24), 168 for 9 (CACC <= 50).
template<int CACC, int MAXW, int CU>
__global__ void __launch_bounds__(32 * (MAXW + 1), 1)
lanczos_step_kernel(const __grid_constant__ StepMaps maps, StepLayout L, const double *__restrict__ coef,
                    double *__restrict__ wi, int have_prev, double *__restrict__ wnext, int qcol,
                    const double *__restrict__ alpha_dev, const double *__restrict__ wsq_dev,
                    long long rows, int TR, int stages, int cgroups, double *__restrict__ partial)
{
    using namespace rails::tma;
    extern __shared__ __align__(128) double ls_smem[];
    __shared__ unsigned long long full[LS_MAX_STAGES], empty[LS_MAX_STAGES];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int total = L.total;
    const int nrg = TR / 32;                            // row groups
    const int ncons = nrg * cgroups;                    // consumer warps
    const size_t stage_doubles = (size_t)(L.padded + 2) * TR;
    double *cs = ls_smem + stages * stage_doubles;      // coefficients
    double *wsum = cs + ((total + 1) & ~1);             // per-warp sums at the end
    double *abuf = wsum + LS_WSUM_DOUBLES;              // partial chain sums [2][groups][TR]
    int *coff = reinterpret_cast<int *>(abuf + (cgroups > 1 ? 2 * cgroups * TR : 0));   // column -> offset in a slab
    int *req_dst = coff + total + 1;                    // requests of one row slab: slab offset,
    int *req_map = req_dst + 64;                        //   tensor map,
    int *req_col = req_map + 64;                        //   first column

    // requests of one slab: per block ceil(n / bc) column boxes, then the basis box
    int nreq = 1;
    unsigned slab_bytes = 2u * LS_BOX_ROWS * sizeof(double);
    for (int b = 0; b < L.nseg; ++b)
    {
        const int boxes = (L.n[b] + L.bc[b] - 1) / L.bc[b];
        nreq += boxes;
        slab_bytes += (unsigned)(boxes * L.bc[b] * LS_BOX_ROWS * sizeof(double));
    }
    if (tid == 0)
    {
        int q = 0, j = 0;
        for (int b = 0; b < L.nseg; ++b)
        {
            for (int c0 = 0; c0 < L.n[b]; c0 += L.bc[b], ++q)
            {
                req_dst[q] = (L.off[b] + c0) * LS_BOX_ROWS;
                req_map[q] = b;
                req_col[q] = c0;
            }
            for (int c = 0; c < L.n[b]; ++c, ++j)
                coff[j] = (L.off[b] + c) * LS_BOX_ROWS;
        }
        req_dst[q] = L.padded * LS_BOX_ROWS;
        req_map[q] = MAX_SEGS;
        req_col[q] = qcol - 1;
        for (int s = 0; s < stages; ++s)
        {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], ncons);
        }
        mbar_fence_init();
    }
    for (int i = tid; i < total; i += blockDim.x) cs[i] = coef[i];
    __syncthreads();

    const long long ntiles = (rows + TR - 1) / TR;
    const int row_slabs = TR / LS_BOX_ROWS;
    const size_t slab_doubles = (size_t)(L.padded + 2) * LS_BOX_ROWS;

    if (warp == ncons)
    {
        // ---- producer warp: keeps `stages` tiles requested ahead of the consumers
        int s = 0;
        unsigned round = 0;
        for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x)
        {
            if (round > 0)
                mbar_wait(&empty[s], (round - 1) & 1u);          // every consumer warp has left stage s
            double *dst = ls_smem + s * stage_doubles;
            if (lane == 0)
                mbar_expect_tx(&full[s], slab_bytes * (unsigned)row_slabs);
            __syncwarp();
            for (int q = lane; q < nreq * row_slabs; q += 32)
            {
                const int rb = q / nreq, e = q - rb * nreq;
                tensor2d_g2s(dst + rb * slab_doubles + req_dst[e], &maps.m[req_map[e]],
                             (int)(tile * TR + (long long)rb * LS_BOX_ROWS), req_col[e], &full[s]);
            }
            if (++s == stages) { s = 0; ++round; }
        }
        return;
    }
    if (warp > ncons)
        return;

    // ---- consumer warp
    const int rg = warp % nrg, cg = warp / nrg;
    const int c0 = cg * CACC, c1 = min(total, c0 + CACC);             // columns of this warp
    const double nrm = sqrt(fabs(*wsq_dev));
    const double alpha = *alpha_dev, sc = 1.0 / nrm, beta = nrm;
    double acc[CACC + 1];
#pragma unroll
    for (int j = 0; j <= CACC; ++j) acc[j] = 0.0;
    const int rt = rg * 32 + lane;                                     // row inside the tile
    const size_t my = (size_t)(rt / LS_BOX_ROWS) * slab_doubles + (rt % LS_BOX_ROWS);

    int s = 0, par = 0;
    unsigned round = 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x)
    {
        const long long r = tile * TR + rt;
        const double *X = ls_smem + s * stage_doubles + my;            // column offset o of this row: X[o]
        mbar_wait(&full[s], round & 1u);
        // the fma chain in block order (the arithmetic of multi_gemv_kernel), over this
        // warp's columns
        // (CU columns at a time: their loads issue ahead of the dependent fmas; as many as
        // the instance's registers hold without spilling)
        double a = 0.0;
        if (CU > 0)
        {
#pragma unroll 1
            for (int jb = c0; jb < c1; jb += (CU > 0 ? CU : 1))
            {
#pragma unroll
                for (int j = 0; j < CU; ++j)
                    if (jb + j < c1)
                        a = fma(X[coff[jb + j]], cs[jb + j], a);
            }
        }
        else
        {
#pragma unroll 8
            for (int j = c0; j < c1; ++j)
                a = fma(X[coff[j]], cs[j], a);
        }
        if (cgroups > 1)
        {
            // partial sums of the row group's warps, added in group order.  Two buffers:
            // a warp is at most one barrier ahead of the others of its group.
            double *ab = abuf + (size_t)par * cgroups * TR;
            ab[cg * TR + rt] = a;
            asm volatile("bar.sync %0, %1;" ::"r"(2 + rg), "r"(cgroups * 32) : "memory");
            a = ab[rt];
            for (int g = 1; g < cgroups; ++g)
                a += ab[g * TR + rt];
            par ^= 1;
        }
        // the update with separately rounded products (the arithmetic of md_tile<., true>)
        const double q = __dmul_rn(X[(size_t)(L.padded + 1) * LS_BOX_ROWS], sc);
        double x = __dadd_rn(a, -__dmul_rn(alpha, q));
        if (have_prev)
            x = __dadd_rn(x, -__dmul_rn(beta, X[(size_t)L.padded * LS_BOX_ROWS]));
        if (r < rows)
        {
            if (cg == 0)
            {
                if (wi != wnext)            // CGS2 use: the column is updated in place
                    wi[r] = q;
                wnext[r] = x;
            }
        }
        else
            x = 0.0;                                                   // rows past the end stay out of the sums
        // the projection of the new vector on this warp's columns, from the same staged tile
#pragma unroll
        for (int j = 0; j < CACC; ++j)
            if (c0 + j < c1)
                acc[j] = fma(X[coff[c0 + j]], x, acc[j]);
        acc[CACC] = fma(x, x, acc[CACC]);
        __syncwarp();
        if (lane == 0)
            mbar_arrive(&empty[s]);                                    // this warp is done with stage s
        if (++s == stages) { s = 0; ++round; }
    }
    // lanes -> warp sums -> CTA sums, once per kernel, fixed order
#pragma unroll
    for (int j = 0; j <= CACC; ++j)
    {
        double p = acc[j];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
            p += __shfl_xor_sync(0xffffffffu, p, o);
        if (lane == 0)
            wsum[warp * (CACC + 1) + j] = p;
    }
    // consumers only (the producer has returned): named barrier 1
    asm volatile("bar.sync 1, %0;" ::"r"(ncons * 32) : "memory");
    for (int j = tid; j <= total; j += ncons * 32)
    {
        // column j lives in column group j / CACC; the squared norm with group 0
        const int g = j < total ? j / CACC : 0, jj = j < total ? j - g * CACC : CACC;
        double p = 0.0;
        for (int w = 0; w < nrg; ++w)
            p += wsum[(g * nrg + w) * (CACC + 1) + jj];
        partial[(size_t)blockIdx.x * (total + 1) + j] = p;
    }
}

// cuTensorMapEncodeTiled through the runtime (no link-time dependency on libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled()
{
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried)
    {
        tried = true;
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
        else
            cudaGetLastError();
    }
    return fn;
}

// column-major block of `cols` columns of `rows` doubles, leading dimension ld: boxes of
// [box_cols][128 rows]
bool make_map(CUtensorMap *m, const double *base, long long rows, int cols, long long ld, int box_cols)
{
    EncodeTiledFn fn = encode_tiled();
    if (!fn || cols < 1 || rows < 1) return false;
    cuuint64_t gdim[2] = {(cuuint64_t)rows, (cuuint64_t)cols};
    cuuint64_t gstride[1] = {(cuuint64_t)ld * sizeof(double)};
    cuuint32_t box[2] = {(cuuint32_t)LS_BOX_ROWS, (cuuint32_t)box_cols};
    cuuint32_t estride[2] = {1, 1};
    return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<double *>(base), gdim, gstride, box, estride,
              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// Optional timing of the step kernel alone (bench.py's roofline): CUDA events on the
// launching stream around every launch, read back by rails_lanczos_step_timing_read.
struct StepTiming
{
    bool on = false;
    std::vector<cudaEvent_t> ev;       // pairs
    size_t used = 0;
    int cols = 0, tile_rows = 0, stages = 0, groups = 0, ctas = 0;
};
StepTiming g_step_timing;

// launches the instance whose per-warp accumulators cover the columns of one group
int launch_lanczos_step_untimed(int grid, size_t smem, StepMaps const &maps, StepLayout const &L, const double *coef,
                                double *wi, int have_prev, double *wnext, int qcol, const double *alpha_dev,
                                const double *wsq_dev, long long rows, int TR, int stages, double *partial);

int launch_lanczos_step(int grid, size_t smem, StepMaps const &maps, StepLayout const &L, const double *coef,
                        double *wi, int have_prev, double *wnext, int qcol, const double *alpha_dev,
                        const double *wsq_dev, long long rows, int TR, int stages, double *partial)
{
    StepTiming &t = g_step_timing;
    if (!t.on)
        return launch_lanczos_step_untimed(grid, smem, maps, L, coef, wi, have_prev, wnext, qcol, alpha_dev, wsq_dev,
                                           rows, TR, stages, partial);
    if (t.used + 2 > t.ev.size())
        for (int i = 0; i < 2; ++i)
        {
            cudaEvent_t e;
            RAILS_CUDA_TRY(cudaEventCreate(&e));
            t.ev.push_back(e);
        }
    RAILS_CUDA_TRY(cudaEventRecord(t.ev[t.used], ctx().stream));
    const int rc = launch_lanczos_step_untimed(grid, smem, maps, L, coef, wi, have_prev, wnext, qcol, alpha_dev,
                                               wsq_dev, rows, TR, stages, partial);
    RAILS_CUDA_TRY(cudaEventRecord(t.ev[t.used + 1], ctx().stream));
    t.used += 2;
    t.cols = L.total; t.tile_rows = TR; t.stages = stages; t.groups = ls_groups(L.total); t.ctas = grid;
    return rc;
}

int launch_lanczos_step_untimed(int grid, size_t smem, StepMaps const &maps, StepLayout const &L, const double *coef,
                        double *wi, int have_prev, double *wnext, int qcol, const double *alpha_dev,
                        const double *wsq_dev, long long rows, int TR, int stages, double *partial)
{
    static int configured_device = -1;
    if (configured_device != ctx().device)
    {
        const int bytes = (int)(LS_SMEM_BUDGET);
        RAILS_CUDA_TRY(cudaFuncSetAttribute(lanczos_step_kernel<12, 16, 12>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
        RAILS_CUDA_TRY(cudaFuncSetAttribute(lanczos_step_kernel<24, 16, 12>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
        RAILS_CUDA_TRY(cudaFuncSetAttribute(lanczos_step_kernel<36, 8, 36>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
        RAILS_CUDA_TRY(cudaFuncSetAttribute(lanczos_step_kernel<50, 8, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
        configured_device = ctx().device;
    }
    const int groups = ls_groups(L.total);
    const int threads = 32 * (TR / 32 * groups + 1);
#define RAILS_LS_ARGS maps, L, coef, wi, have_prev, wnext, qcol, alpha_dev, wsq_dev, rows, TR, stages, groups, partial
    if (groups > 1)
        RAILS_LAUNCH((lanczos_step_kernel<24, 16, 12>), grid, threads, smem, RAILS_LS_ARGS);
    else if (L.total <= 12)
        RAILS_LAUNCH((lanczos_step_kernel<12, 16, 12>), grid, threads, smem, RAILS_LS_ARGS);
    else if (L.total <= 24)
        RAILS_LAUNCH((lanczos_step_kernel<24, 16, 12>), grid, threads, smem, RAILS_LS_ARGS);
    else if (L.total <= 36)
        RAILS_LAUNCH((lanczos_step_kernel<36, 8, 36>), grid, threads, smem, RAILS_LS_ARGS);
    else
    {
        RAILS_LAUNCH((lanczos_step_kernel<50, 8, 0>), grid, threads, smem, RAILS_LS_ARGS);
    }
#undef RAILS_LS_ARGS
    RAILS_CHECK_LAUNCH();
    return RAILS_OK;
}

__device__ double g_cgs2_consts[2] = {-1.0, 1.0};      // alpha = -1, |w|^2 = 1 for the CGS2 use of the step kernel

__global__ void __launch_bounds__(128)
negate_kernel(const double *__restrict__ in, double *__restrict__ out, int n)
{
    for (int i = threadIdx.x; i < n; i += 128) out[i] = -in[i];
}

// q = w / |w| for the last column of the recurrence
__global__ void __launch_bounds__(256)
normalize_kernel(double *__restrict__ w, const double *__restrict__ sq_dev, long long rows)
{
    const double s = 1.0 / sqrt(fabs(*sq_dev));
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < rows; i += (long long)gridDim.x * 256)
        w[i] = __dmul_rn(w[i], s);
}

// out[i] = sum_b partial[b * n + i], one CTA per output, fixed order:
// thread t adds blocks t, t+256, ...; then a shared-memory tree.
__global__ void __launch_bounds__(256)
sum_partials_kernel(const double *__restrict__ partial, int nblocks, int n,
                    double *__restrict__ out, int take_sqrt)
{
    __shared__ double red[256];
    const int i = blockIdx.x;
    double s = 0.0;
    for (int b = threadIdx.x; b < nblocks; b += 256)
        s += partial[(size_t)b * n + i];
    red[threadIdx.x] = s;
    __syncthreads();
    for (int w = 128; w > 0; w >>= 1)
    {
        if (threadIdx.x < w)
            red[threadIdx.x] += red[threadIdx.x + w];
        __syncthreads();
    }
    if (threadIdx.x == 0)
        out[i] = take_sqrt ? sqrt(fabs(red[0])) : red[0];
}

int collect_segs(int nblk, const rails_mv_t *X, const int *xc0, const int *xn, long long rows,
                 Segs *S, int *total, bool *distributed, const char *what)
{
    if (nblk < 1 || nblk > MAX_SEGS || !X || !xc0 || !xn)
        return rails::fail(RAILS_ERR_ARG, "%s: between 1 and %d column blocks", what, MAX_SEGS);
    S->nseg = 0;
    *total = 0;
    *distributed = false;
    for (int b = 0; b < nblk; ++b)
    {
        if (!X[b] || xc0[b] < 0 || xn[b] < 0 || xc0[b] + xn[b] > X[b]->cols)
            return rails::fail(RAILS_ERR_ARG, "%s: block %d column range outside capacity", what, b);
        if (X[b]->rows != rows)
            return rails::fail(RAILS_ERR_ARG, "%s: block %d has %lld rows, expected %lld", what, b,
                               X[b]->rows, rows);
        if (X[b]->flags == RAILS_MV_DISTRIBUTED)
            *distributed = true;
        *total += xn[b];
        if (xn[b] == 0)
            continue;                 // empty blocks keep their (empty) slot in the output
        const int s = S->nseg++;
        S->p[s] = X[b]->data + (size_t)xc0[b] * X[b]->ld;
        S->ld[s] = X[b]->ld;
        S->n[s] = xn[b];
    }
    return RAILS_OK;
}

// device result (n doubles at c.d_small) -> all-reduce over ranks -> host
int finish_to_host(int n, bool distributed, double *out)
{
    rails::Context &c = ctx();
    if (distributed && c.comm)
    {
        int rc = rails::finish_reduction(c.d_small, 1, n, c.d_small, true);
        if (rc) return rc;
    }
    RAILS_CUDA_TRY(cudaMemcpyAsync(c.h_small, c.d_small, (size_t)n * sizeof(double),
                                   cudaMemcpyDeviceToHost, c.stream));
    RAILS_CUDA_TRY(cudaStreamSynchronize(c.stream));
    { int xrc = rails::check_exchange_error("reduction"); if (xrc) return xrc; }
    memcpy(out, c.h_small, (size_t)n * sizeof(double));
    return RAILS_OK;
}

}

namespace rails
{

int scal_inv_norm_device(double *x, long long ld, long long rows, const double *d);

// CGS2 of one column with the step kernel (round 2): the column v = V[:, vcol] against the
// `against` columns before it, c1 = Vo'v already on the device.  The reference's two
// passes (StlWrapper::orthogonalize, src/StlWrapper.cpp:305-321: dot, update, dot,
// update, norm) read Vo four times; here
//     step kernel:  v <- v - Vo c1,  c2 = Vo'v and |v|^2 from the same staged tile
//     step kernel:  v <- v - Vo c2,  |v|^2                 (its Vo'v is not used)
//     v <- v / |v|
// read it twice more after the first dot: three passes instead of four, the same sums
// (v - Vo c is formed as fma chain + v, exactly the update of gemm_kernel).
// Returns 1 when the shape does not fit the step kernel (the caller runs the plain chain).
int cgs2_fused(rails_mv_s *V, int first_col, int against, int vcol, const double *c1_dev, bool dist)
{
    Context &c = ctx();
    // RAILS_CGS2_FUSED=0, or RAILS_FUSED_LANCZOS=0 (the operation-by-operation mode, which
    // keeps the reference's call sequence everywhere), select the plain chain; read per
    // call, the tests switch modes inside one process
    const char *e1 = getenv("RAILS_CGS2_FUSED"), *e2 = getenv("RAILS_FUSED_LANCZOS");
    if ((e1 && e1[0] == '0') || (e2 && strcmp(e2, "0") == 0) || against < 1 || vcol < 1)
        return 1;
    int n1[1] = {against};
    StepLayout lay = ls_layout(n1, 1);
    int tile_rows = 0, stages = 0;
    ls_shape(lay, &tile_rows, &stages);
    if (tile_rows == 0 || !encode_tiled())
        return 1;
    const long long rows = V->rows;
    static double *scratch = nullptr, *consts = nullptr;
    static int scratch_device = -1;
    if (scratch_device != c.device)
    {
        int rc = dev_alloc_t(&scratch, (size_t)2 * (LS_MAX_COLS + 2) * sizeof(double));
        if (rc) return rc;
        RAILS_CUDA_TRY(cudaGetSymbolAddress((void **)&consts, g_cgs2_consts));
        scratch_device = c.device;
    }
    double *coef = scratch, *g = scratch + LS_MAX_COLS + 2;
    StepMaps maps;
    bool have = rows > 0;
    if (have)
        have = make_map(&maps.m[0], V->data + (size_t)first_col * V->ld, rows, against, V->ld, lay.bc[0]) &&
               make_map(&maps.m[MAX_SEGS], V->data, rows, V->cols, V->ld, 2);
    if (rows > 0 && !have)
        return 1;
    const long long tiles = (rows + tile_rows - 1) / tile_rows;
    const int nb = (int)std::max<long long>(1, std::min<long long>(tiles, c.sm_count));
    int rc = ensure_partial((size_t)nb * (against + 1));
    if (rc) return rc;
    double *v = V->data + (size_t)vcol * V->ld;
    const double *cin = c1_dev;
    for (int pass = 0; pass < 2; ++pass)
    {
        RAILS_LAUNCH(negate_kernel, 1, 128, 0, cin, coef, against);
        if (rows > 0)
        {
            rc = launch_lanczos_step(nb, ls_smem_bytes(lay, tile_rows, stages), maps, lay, (const double *)coef, v, 0, v,
                                     vcol, (const double *)consts, (const double *)(consts + 1), rows, tile_rows, stages,
                                     c.d_partial);
            if (rc) return rc;
        }
        else
            RAILS_CUDA_TRY(cudaMemsetAsync(c.d_partial, 0, (size_t)nb * (against + 1) * sizeof(double), c.stream));
        rc = finish_reduction(c.d_partial, nb, against + 1, g, dist, nullptr, 0, nullptr, nullptr, nullptr, nullptr);
        if (rc) return rc;
        cin = g;
    }
    return scal_inv_norm_device(v, V->ld, rows, g + against);
}

}

// out (host, sum of xn doubles) = [X_0 X_1 ...]' (qscale * q); when
// qscale != 1 the scaled q is written back in the same pass.
extern "C" int rails_mv_multi_dot(int nblk, const rails_mv_t *X, const int *xc0, const int *xn,
                                  rails_mv_t q, int qc, double qscale, double *out)
{
    RAILS_REQUIRE_READY();
    if (!q || qc < 0 || qc >= q->cols || !out)
        return rails::fail(RAILS_ERR_ARG, "multi_dot: bad q or result");
    Segs S;
    int total = 0;
    bool dist = false;
    int rc = collect_segs(nblk, X, xc0, xn, q->rows, &S, &total, &dist, "multi_dot");
    if (rc) return rc;
    if (total > MD_MAXCOLS)
        return rails::fail(RAILS_ERR_ARG, "multi_dot: %d columns, at most %d per call", total, MD_MAXCOLS);
    for (int s = 0; s < S.nseg; ++s)
        if (S.p[s] <= q->data + (size_t)qc * q->ld && q->data + (size_t)qc * q->ld < S.p[s] + (size_t)S.n[s] * S.ld[s]
            && qscale != 1.0)
            return rails::fail(RAILS_ERR_ARG, "multi_dot: q is rescaled and lies inside a block");
    rails::Context &c = ctx();
    const long long rows = q->rows;
    const int write_q = qscale != 1.0 ? 1 : 0;
    if (total == 0 && !write_q)
        return RAILS_OK;
    rc = rails::ensure_small((size_t)std::max(total, 1));
    if (rc) return rc;
    if (rows == 0)
    {
        RAILS_CUDA_TRY(cudaMemsetAsync(c.d_small, 0, (size_t)std::max(total, 1) * sizeof(double), c.stream));
    }
    else
    {
        long long ntiles = (rows + MD_TILE - 1) / MD_TILE;
        const int nb = (int)std::min<long long>(ntiles, (long long)c.sm_count * 4);
        rc = rails::ensure_partial((size_t)nb * std::max(total, 1));
        if (rc) return rc;
        static int configured_device = -1;
        if (configured_device != c.device)
        {
            RAILS_CUDA_TRY(cudaFuncSetAttribute(multi_dot_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                MD_MAXCOLS * MD_WARPS * (int)sizeof(double)));
            configured_device = c.device;
        }
        const size_t smem = (size_t)std::max(total, 1) * MD_WARPS * sizeof(double);
        RAILS_LAUNCH(multi_dot_kernel, nb, MD_THREADS, smem, S, total, q->data + (size_t)qc * q->ld,
                     qscale, write_q, rows, c.d_partial, (const double *)nullptr);
        RAILS_CHECK_LAUNCH();
        if (total > 0)
        {
            RAILS_LAUNCH(sum_partials_kernel, total, 256, 0, c.d_partial, nb, total, c.d_small, 0);
            RAILS_CHECK_LAUNCH();
        }
    }
    if (total == 0)
        return RAILS_OK;
    return finish_to_host(total, dist && q->flags == RAILS_MV_DISTRIBUTED, out);
}

namespace
{

constexpr int MG_THREADS = 256;
constexpr int MG_RP = 2;              // double2 per thread: 4 rows
constexpr int MG_U = 8;               // columns in flight

// out = sum_s X_s coef_s  (one output column), optionally partial[blockIdx.x]
// = sum over this CTA's rows of out * q.  The sum over all columns is one
// fma chain per row, in block order.
__global__ void __launch_bounds__(MG_THREADS)
multi_gemv_kernel(Segs S, int total, const double *__restrict__ coef, double *__restrict__ out,
                  const double *__restrict__ q, long long rows, double *__restrict__ partial)
{
    const long long rows2 = (rows + 1) / 2;
    extern __shared__ double cs[];
    __shared__ double red[MG_THREADS / 32];
    for (int i = threadIdx.x; i < total; i += MG_THREADS)
        cs[i] = coef[i];
    __syncthreads();

    long long r2[MG_RP];
#pragma unroll
    for (int v = 0; v < MG_RP; ++v)
        r2[v] = ((long long)blockIdx.x * MG_RP + v) * MG_THREADS + threadIdx.x;
    double2 acc[MG_RP];
#pragma unroll
    for (int v = 0; v < MG_RP; ++v)
        acc[v] = make_double2(0.0, 0.0);

    int off = 0;
#pragma unroll
    for (int s = 0; s < MAX_SEGS; ++s)
    {
        if (s >= S.nseg) break;
        const double2 *__restrict__ X = reinterpret_cast<const double2 *>(S.p[s]);
        const long long ld2 = S.ld[s] / 2;
        const int kc = S.n[s];
        int j = 0;
        for (; j + MG_U <= kc; j += MG_U)
        {
            double2 x[MG_U][MG_RP];
#pragma unroll
            for (int u = 0; u < MG_U; ++u)
#pragma unroll
                for (int v = 0; v < MG_RP; ++v)
                    x[u][v] = (r2[v] < rows2) ? __ldg(X + (size_t)(j + u) * ld2 + r2[v])
                                              : make_double2(0.0, 0.0);
#pragma unroll
            for (int u = 0; u < MG_U; ++u)
            {
                const double cf = cs[off + j + u];
#pragma unroll
                for (int v = 0; v < MG_RP; ++v)
                {
                    acc[v].x = fma(x[u][v].x, cf, acc[v].x);
                    acc[v].y = fma(x[u][v].y, cf, acc[v].y);
                }
            }
        }
        for (; j < kc; ++j)
        {
            const double cf = cs[off + j];
#pragma unroll
            for (int v = 0; v < MG_RP; ++v)
            {
                const double2 x = (r2[v] < rows2) ? __ldg(X + (size_t)j * ld2 + r2[v])
                                                  : make_double2(0.0, 0.0);
                acc[v].x = fma(x.x, cf, acc[v].x);
                acc[v].y = fma(x.y, cf, acc[v].y);
            }
        }
        off += kc;
    }

    double dotp = 0.0;
#pragma unroll
    for (int v = 0; v < MG_RP; ++v)
    {
        if (r2[v] >= rows2) continue;
        reinterpret_cast<double2 *>(out)[r2[v]] = acc[v];
        if (q)
        {
            const double2 qq = __ldg(reinterpret_cast<const double2 *>(q) + r2[v]);
            dotp = fma(acc[v].x, qq.x, dotp);
            if (2 * r2[v] + 1 < rows)          // the pad element of an odd row count stays out
                dotp = fma(acc[v].y, qq.y, dotp);
        }
    }
    if (q)
    {
        for (int o = 16; o > 0; o >>= 1)
            dotp += __shfl_xor_sync(0xffffffffu, dotp, o);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = dotp;
        __syncthreads();
        if (threadIdx.x == 0)
        {
            double t = 0.0;
#pragma unroll
            for (int w = 0; w < MG_THREADS / 32; ++w) t += red[w];
            partial[blockIdx.x] = t;
        }
    }
}

constexpr int LU_THREADS = 256;

// next = (next - round(alpha q)) - round(beta p)   (p optional), and
// partial[blockIdx.x] = sum of next^2 over this CTA's rows (:411-417)
__global__ void __launch_bounds__(LU_THREADS)
lanczos_update_kernel(double *__restrict__ next, const double *__restrict__ q,
                      const double *__restrict__ p, double alpha, double beta, long long rows,
                      double *__restrict__ partial, const double *__restrict__ alpha_dev,
                      const double *__restrict__ beta_sq_dev)
{
    __shared__ double red[LU_THREADS / 32];
    // device-resident recurrence: alpha and |next_prev|^2 come from device memory
    if (alpha_dev) alpha = *alpha_dev;
    if (beta_sq_dev) beta = sqrt(fabs(*beta_sq_dev));
    const long long rows2 = rows / 2;
    double2 *n2 = reinterpret_cast<double2 *>(next);
    const double2 *q2 = reinterpret_cast<const double2 *>(q);
    const double2 *p2 = reinterpret_cast<const double2 *>(p);
    double s = 0.0;
    for (long long i = (long long)blockIdx.x * LU_THREADS + threadIdx.x; i < rows2;
         i += (long long)gridDim.x * LU_THREADS)
    {
        double2 x = n2[i];
        const double2 a = __ldg(q2 + i);
        x.x = __dadd_rn(x.x, -__dmul_rn(alpha, a.x));
        x.y = __dadd_rn(x.y, -__dmul_rn(alpha, a.y));
        if (p)
        {
            const double2 b = __ldg(p2 + i);
            x.x = __dadd_rn(x.x, -__dmul_rn(beta, b.x));
            x.y = __dadd_rn(x.y, -__dmul_rn(beta, b.y));
        }
        n2[i] = x;
        s = fma(x.x, x.x, s);
        s = fma(x.y, x.y, s);
    }
    if ((rows & 1) && blockIdx.x == 0 && threadIdx.x == 0)
    {
        const long long r = rows - 1;
        double x = __dadd_rn(next[r], -__dmul_rn(alpha, q[r]));
        if (p) x = __dadd_rn(x, -__dmul_rn(beta, p[r]));
        next[r] = x;
        s = fma(x, x, s);
    }
    for (int o = 16; o > 0; o >>= 1)
        s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0)
    {
        double t = 0.0;
#pragma unroll
        for (int w = 0; w < LU_THREADS / 32; ++w) t += red[w];
        partial[blockIdx.x] = t;
    }
}

}

// out[:, oc] = sum_b X_b[:, xc0_b : xc0_b + xn_b) * coef_b, coef = the blocks'
// coefficient vectors concatenated (host).  With q != NULL also
// *alpha = out[:, oc]' q[:, qc] (summed over ranks).
extern "C" int rails_mv_multi_gemv(int nblk, const rails_mv_t *X, const int *xc0, const int *xn,
                                   const double *coef, rails_mv_t out, int oc,
                                   rails_mv_t q, int qc, double *alpha)
{
    RAILS_REQUIRE_READY();
    if (!out || oc < 0 || oc >= out->cols)
        return rails::fail(RAILS_ERR_ARG, "multi_gemv: bad result column");
    if (q && (qc < 0 || qc >= q->cols || q->rows != out->rows || !alpha))
        return rails::fail(RAILS_ERR_ARG, "multi_gemv: bad q");
    Segs S;
    int total = 0;
    bool dist = false;
    int rc = collect_segs(nblk, X, xc0, xn, out->rows, &S, &total, &dist, "multi_gemv");
    if (rc) return rc;
    if (total > MD_MAXCOLS)
        return rails::fail(RAILS_ERR_ARG, "multi_gemv: %d columns, at most %d per call", total, MD_MAXCOLS);
    if (total > 0 && !coef)
        return rails::fail(RAILS_ERR_ARG, "multi_gemv: null coefficients");
    double *o = out->data + (size_t)oc * out->ld;
    for (int s = 0; s < S.nseg; ++s)
        if (S.p[s] <= o && o < S.p[s] + (size_t)S.n[s] * S.ld[s])
            return rails::fail(RAILS_ERR_ARG, "multi_gemv: result column lies inside a block");
    rails::Context &c = ctx();
    const long long rows = out->rows;
    if (alpha) *alpha = 0.0;
    // (a rank without rows still launches: it has to take part in the sum over ranks)
    const double *dC = nullptr;
    if (total > 0)
    {
        rc = rails::upload_small(coef, total, 1, total, &dC);
        if (rc) return rc;
    }
    const long long rows2 = (rows + 1) / 2;
    const long long per_block = (long long)MG_THREADS * MG_RP;
    const int nb = (int)std::max<long long>(1, (rows2 + per_block - 1) / per_block);
    rc = rails::ensure_partial((size_t)nb);
    if (rc) return rc;
    rc = rails::ensure_small(1);
    if (rc) return rc;
    const size_t smem = (size_t)std::max(total, 1) * sizeof(double);
    RAILS_LAUNCH(multi_gemv_kernel, nb, MG_THREADS, smem, S, total, dC, o,
                 q ? q->data + (size_t)qc * q->ld : (const double *)nullptr, rows, c.d_partial);
    RAILS_CHECK_LAUNCH();
    if (!q)
        return RAILS_OK;
    RAILS_LAUNCH(sum_partials_kernel, 1, 256, 0, c.d_partial, nb, 1, c.d_small, 0);
    RAILS_CHECK_LAUNCH();
    return finish_to_host(1, out->flags == RAILS_MV_DISTRIBUTED, alpha);
}

// next[:, nc] -= alpha * q[:, qc];  next[:, nc] -= beta * p[:, pc] (p may be
// NULL);  *norm = |next[:, nc]|_2 over all ranks
// (src/LyapunovSolver.hpp:411-417 in one pass)
extern "C" int rails_mv_lanczos_update(rails_mv_t next, int nc, double alpha, rails_mv_t q, int qc,
                                       double beta, rails_mv_t p, int pc, double *norm)
{
    RAILS_REQUIRE_READY();
    if (!next || !q || !norm || nc < 0 || nc >= next->cols || qc < 0 || qc >= q->cols ||
        q->rows != next->rows)
        return rails::fail(RAILS_ERR_ARG, "lanczos_update: bad next or q");
    if (p && (pc < 0 || pc >= p->cols || p->rows != next->rows))
        return rails::fail(RAILS_ERR_ARG, "lanczos_update: bad p");
    rails::Context &c = ctx();
    const long long rows = next->rows;
    *norm = 0.0;
    int rc = rails::ensure_small(1);
    if (rc) return rc;
    if (rows == 0)
    {
        RAILS_CUDA_TRY(cudaMemsetAsync(c.d_small, 0, sizeof(double), c.stream));
    }
    else
    {
        long long want = (rows / 2 + LU_THREADS - 1) / LU_THREADS;
        const int nb = (int)std::max<long long>(1, std::min<long long>(want, (long long)c.sm_count * 8));
        rc = rails::ensure_partial((size_t)nb);
        if (rc) return rc;
        RAILS_LAUNCH(lanczos_update_kernel, nb, LU_THREADS, 0, next->data + (size_t)nc * next->ld,
                     q->data + (size_t)qc * q->ld,
                     p ? p->data + (size_t)pc * p->ld : (const double *)nullptr, alpha, beta, rows,
                     c.d_partial, (const double *)nullptr, (const double *)nullptr);
        RAILS_CHECK_LAUNCH();
        RAILS_LAUNCH(sum_partials_kernel, 1, 256, 0, c.d_partial, nb, 1, c.d_small, 0);
        RAILS_CHECK_LAUNCH();
    }
    double sq = 0.0;
    rc = finish_to_host(1, next->flags == RAILS_MV_DISTRIBUTED, &sq);
    if (rc) return rc;
    *norm = std::sqrt(std::fabs(sq));
    return RAILS_OK;
}

// Kernel A of the device-resident recurrence on its own (unit tests, bench.py's
// roofline leg): with prev_col >= 0 column `col` of Q is first replaced by
//     w = Q[:, col] - alpha * (Q[:, prev_col] / |Q[:, prev_col]|) - |Q[:, prev_col]| * Q[:, prev2_col]
// (prev2_col < 0: no third term) and Q[:, prev_col] is normalised in place; then
// out[0 .. total) = [X_0 X_1 ..]' w and out[total] = |w|^2 (summed over ranks).
extern "C" int rails_mv_lanczos_project(int nblk, const rails_mv_t *X, const int *xc0, const int *xn,
                                        rails_mv_t Q, int col, int prev_col, int prev2_col,
                                        double alpha, double *out)
{
    RAILS_REQUIRE_READY();
    if (!Q || !out || col < 0 || col >= Q->cols || prev_col >= Q->cols || prev2_col >= Q->cols ||
        (prev_col < 0 && prev2_col >= 0) || prev_col == col || prev2_col == col)
        return rails::fail(RAILS_ERR_ARG, "lanczos_project: bad columns");
    Segs S;
    int total = 0;
    bool dist = false;
    int rc = collect_segs(nblk, X, xc0, xn, Q->rows, &S, &total, &dist, "lanczos_project");
    if (rc) return rc;
    if (total > MD_MAXCOLS)
        return rails::fail(RAILS_ERR_ARG, "lanczos_project: %d columns, at most %d", total, MD_MAXCOLS);
    rails::Context &c = ctx();
    const long long rows = Q->rows;
    const bool reduce = dist && Q->flags == RAILS_MV_DISTRIBUTED && c.comm;
    rc = rails::ensure_small((size_t)total + 4);
    if (rc) return rc;
    double *g = c.d_small, *sc = g + total + 1;        // sc[0] = alpha, sc[1] = |prev|^2
    double *colp = Q->data + (size_t)col * Q->ld;
    double *prev = prev_col >= 0 ? Q->data + (size_t)prev_col * Q->ld : nullptr;
    const double *prev2 = prev2_col >= 0 ? Q->data + (size_t)prev2_col * Q->ld : nullptr;
    if (prev)
    {
        const double *ap = nullptr;
        rc = rails::upload_small(&alpha, 1, 1, 1, &ap);
        if (rc) return rc;
        RAILS_CUDA_TRY(cudaMemcpyAsync(sc, ap, sizeof(double), cudaMemcpyDeviceToDevice, c.stream));
        rc = rails::dot_device(prev, Q->ld, 1, prev, Q->ld, 1, rows, sc + 1, 1, reduce);
        if (rc) return rc;
    }
    const long long ntiles = (rows + MD_TILE - 1) / MD_TILE;
    const int nb = (int)std::max<long long>(1, std::min<long long>(ntiles, (long long)c.sm_count * 4));
    rc = rails::ensure_partial((size_t)nb * (total + 1));
    if (rc) return rc;
    rc = lz_attr();
    if (rc) return rc;
    RAILS_LAUNCH(lanczos_dot_kernel, nb, MD_THREADS, (size_t)(total + 1) * MD_WARPS * sizeof(double), S, total,
                 colp, prev, prev2, (const double *)sc, (const double *)(sc + 1), rows, c.d_partial);
    RAILS_CHECK_LAUNCH();
    rc = rails::finish_reduction(c.d_partial, nb, total + 1, g, reduce);
    if (rc) return rc;
    RAILS_CUDA_TRY(cudaMemcpyAsync(c.h_small, g, (size_t)(total + 1) * sizeof(double), cudaMemcpyDeviceToHost, c.stream));
    RAILS_CUDA_TRY(cudaStreamSynchronize(c.stream));
    { int xrc = rails::check_exchange_error("lanczos_project"); if (xrc) return xrc; }
    memcpy(out, c.h_small, (size_t)(total + 1) * sizeof(double));
    return RAILS_OK;
}

namespace
{

constexpr int LZ_MAXL = 256;           // Lanczos steps per call

int lz_attr()
{
    static int configured_device = -1;
    if (configured_device != ctx().device)
    {
        RAILS_CUDA_TRY(cudaFuncSetAttribute(lanczos_dot_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (MD_MAXCOLS + 1) * MD_WARPS * (int)sizeof(double)));
        configured_device = ctx().device;
    }
    return RAILS_OK;
}

int lz_scratch(int total)
{
    rails::Context &c = ctx();
    const size_t need = 2 * (size_t)LZ_MAXL + 1 + 2 * ((size_t)MD_MAXCOLS + 2);
    (void)total;
    if (c.lz_cap >= need) return RAILS_OK;
    RAILS_CUDA_TRY(cudaStreamSynchronize(c.stream));
    if (c.d_lz) RAILS_CUDA_TRY(cudaFree(c.d_lz));
    RAILS_CUDA_TRY(cudaMalloc(&c.d_lz, need * sizeof(double)));
    RAILS_CUDA_TRY(cudaMemset(c.d_lz, 0, need * sizeof(double)));
    c.lz_cap = need;
    return RAILS_OK;
}

}

// The whole L-step recurrence of Solver::resid_lanczos
// (src/LyapunovSolver.hpp:375-434) as ONE asynchronous chain on the library
// stream: no value returns to the host until the last step has been enqueued.
// Blocks 0 and 1 (k columns each: V or MV, and AV) exchange their projections
// through T; the remaining blocks (B) use theirs directly:
//     R q = X_0 (T X_1'q) + X_1 (T X_0'q) + X_2 (X_2'q).
// Q[:, qc0] holds the un-normalised start vector; on return Q[:, qc0 .. qc0+steps)
// are the Lanczos vectors, alpha[i] = H(i,i), beta[i] = H(i+1,i).  A breakdown
// (beta_i < 1e-14, :419-427) is detected afterwards: *steps = i + 1 and what was
// computed past it is ignored, which is what the reference's `break` leaves.
extern "C" int rails_mv_lanczos_run(int nblk, const rails_mv_t *X, const int *xc0, const int *xn,
                                    const double *T, int ldt, rails_mv_t Q, int qc0, int L,
                                    double *alpha, double *beta, int *steps)
{
    RAILS_REQUIRE_READY();
    if (!Q || qc0 < 0 || L < 1 || qc0 + L + 1 > Q->cols || !alpha || !beta || !steps)
        return rails::fail(RAILS_ERR_ARG, "lanczos_run: Q needs L + 1 columns from qc0");
    if (nblk < 2 || xn[0] != xn[1] || (xn[0] > 0 && (!T || ldt < xn[0])))
        return rails::fail(RAILS_ERR_ARG, "lanczos_run: blocks 0 and 1 must have k columns each and T be k x k");
    Segs S;
    int total = 0;
    bool dist = false;
    int rc = collect_segs(nblk, X, xc0, xn, Q->rows, &S, &total, &dist, "lanczos_run");
    if (rc) return rc;
    if (total > MD_MAXCOLS)
        return rails::fail(RAILS_ERR_ARG, "lanczos_run: %d columns, at most %d", total, MD_MAXCOLS);
    for (int s = 0; s < S.nseg; ++s)
        if (S.p[s] < Q->data + (size_t)(qc0 + L + 1) * Q->ld && Q->data + (size_t)qc0 * Q->ld < S.p[s] + (size_t)S.n[s] * S.ld[s])
            return rails::fail(RAILS_ERR_ARG, "lanczos_run: Q overlaps a block");
    rails::Context &c = ctx();
    const int k = xn[0];
    const long long rows = Q->rows;
    const bool reduce = dist && Q->flags == RAILS_MV_DISTRIBUTED && c.comm;
    *steps = 0;

    // device scratch (own buffer: it must survive the calls between
    // rails_mv_lanczos_begin and this one):
    //   b2[LZ_MAXL + 1] | a[LZ_MAXL] | g[total + 1] | coef[total + 1]      (b2[i] = |w_i|^2)
    if (L > LZ_MAXL)
        return rails::fail(RAILS_ERR_ARG, "lanczos_run: at most %d steps", LZ_MAXL);
    const bool started = c.lz_pending && c.lz_q == Q->data + (size_t)qc0 * Q->ld && c.lz_total == total;
    c.lz_pending = false;
    rc = lz_scratch(total);
    if (rc) return rc;
    const int tot1 = std::max(total, 1) + 1;
    double *b2 = c.d_lz, *a = b2 + LZ_MAXL + 1, *g = a + LZ_MAXL, *coef = g + tot1;
    const double *dT = nullptr;
    if (k > 0)
    {
        rc = rails::upload_small(T, k, k, ldt, &dT);
        if (rc) return rc;
    }
    // b2[0] (and g) are already there when the first projection was started early
    RAILS_CUDA_TRY(cudaMemsetAsync(b2 + 1, 0, (2 * (size_t)LZ_MAXL) * sizeof(double), c.stream));
    if (!started)
        RAILS_CUDA_TRY(cudaMemsetAsync(b2, 0, sizeof(double), c.stream));

    const long long ntiles = (rows + MD_TILE - 1) / MD_TILE;
    const int nb_dot = (int)std::max<long long>(1, std::min<long long>(ntiles, (long long)c.sm_count * 4));
    const long long rows2 = (rows + 1) / 2;
    const int nb_gemv = (int)std::max<long long>(1, (rows2 + (long long)MG_THREADS * MG_RP - 1) / ((long long)MG_THREADS * MG_RP));
    rc = rails::ensure_partial(std::max<size_t>((size_t)nb_dot * (total + 1), (size_t)nb_gemv));
    if (rc) return rc;
    rc = lz_attr();
    if (rc) return rc;
    {
        // (a rank that owns no rows still runs every launch, with idle CTAs: it has
        // to take part in each exchange)
        // The Lanczos vectors are kept UN-normalised while the recurrence runs
        // (w_i = |w_i| q_i): a step is two kernels and two exchanges,
        //   A_i: w_i = raw_i - alpha_{i-1} q_{i-1} - beta_{i-2} q_{i-2} (fused update of the
        //        previous step, which also normalises column i-1 in place),
        //        [X_0 X_1 ..]' w_i and |w_i|^2           -> exchange -> g = ./|w_i|, coef
        //   B_i: raw_{i+1} = sum_b X_b coef_b,  raw_{i+1}' w_i -> exchange -> alpha_i = ./|w_i|
        // instead of three (dot, gemv, update) with three.  The update of the
        // last step only feeds H(L, L-1), which the caller drops: it is skipped.
        // alpha_i = q_i' R q_i: by default from the projections the first finish already
        // holds, alpha = g . coef (= sum_b (X_b'q) . coef_b, the same number as
        // raw_{i+1}' q_i in exact arithmetic) -- ONE exchange and three launches per
        // step.  RAILS_LANCZOS_ALPHA=dot computes it as the reference does (:406), as a
        // second reduction over the rows inside kernel B with its own exchange.
        static int alpha_from_dot = -1;
        if (alpha_from_dot < 0)
        {
            const char *e = getenv("RAILS_LANCZOS_ALPHA");
            alpha_from_dot = (e && strcmp(e, "dot") == 0) ? 1 : 0;
        }
        // One kernel per step (lanczos_step_kernel) for 2k + n <= 96 columns, when the
        // driver offers tensor maps.
        // RAILS_LANCZOS_ONE_KERNEL=0 keeps the two-kernel step.
        static int one_kernel = -1;
        if (one_kernel < 0)
        {
            const char *e = getenv("RAILS_LANCZOS_ONE_KERNEL");
            one_kernel = (e && e[0] == '0') ? 0 : 1;
        }
        StepLayout lay = ls_layout(S.n, S.nseg);
        int tile_rows = 0, stages = 0;
        if (total > 0)
            ls_shape(lay, &tile_rows, &stages);
        StepMaps maps;
        bool have_maps = one_kernel && !alpha_from_dot && tile_rows > 0 && L > 1;
        for (int b = 0; have_maps && b < S.nseg; ++b)
            have_maps = make_map(&maps.m[b], S.p[b], rows, S.n[b], S.ld[b], lay.bc[b]);
        if (have_maps)
            have_maps = make_map(&maps.m[MAX_SEGS], Q->data, rows, Q->cols, Q->ld, 2);
        if (have_maps)
        {
            const long long tiles = (rows + tile_rows - 1) / tile_rows;
            const int nb_step = (int)std::max<long long>(1, std::min<long long>(tiles, c.sm_count));
            rc = rails::ensure_partial(std::max<size_t>((size_t)nb_dot * (total + 1), (size_t)nb_step * (total + 1)));
            if (rc) return rc;
            double *col0 = Q->data + (size_t)qc0 * Q->ld;
            if (started)
            {
                rc = rails::lanczos_coef(dT, k, total, g, coef, a);
                if (rc) return rc;
            }
            else
            {
                RAILS_LAUNCH(lanczos_dot_kernel, nb_dot, MD_THREADS, (size_t)(total + 1) * MD_WARPS * sizeof(double),
                             S, total, col0, (double *)nullptr, (const double *)nullptr, (const double *)a,
                             (const double *)b2, rows, c.d_partial);
                RAILS_CHECK_LAUNCH();
                rc = rails::finish_reduction(c.d_partial, nb_dot, total + 1, g, reduce, dT, k, coef, b2, nullptr, a);
                if (rc) return rc;
            }
            for (int i = 0; i + 1 < L; ++i)
            {
                double *col = col0 + (size_t)i * Q->ld;
                rc = launch_lanczos_step(nb_step, ls_smem_bytes(lay, tile_rows, stages), maps, lay, (const double *)coef, col,
                                         i > 0 ? 1 : 0, col + Q->ld, qc0 + i, (const double *)(a + i),
                                         (const double *)(b2 + i), rows, tile_rows, stages, c.d_partial);
                if (rc) return rc;
                rc = rails::finish_reduction(c.d_partial, nb_step, total + 1, g, reduce, dT, k, coef, b2 + i + 1,
                                             nullptr, a + i + 1);
                if (rc) return rc;
            }
        }
        else
        for (int i = 0; i < L; ++i)
        {
            double *col = Q->data + (size_t)(qc0 + i) * Q->ld;
            double *alpha_here = alpha_from_dot ? (double *)nullptr : a + i;
            if (i == 0 && started)
            {
                // A_0 and its exchange ran while the host solved the projected
                // problem (rails_mv_lanczos_begin); only the k x k products remain
                rc = rails::lanczos_coef(dT, k, total, g, coef, alpha_here);
                if (rc) return rc;
            }
            else
            {
                RAILS_LAUNCH(lanczos_dot_kernel, nb_dot, MD_THREADS, (size_t)(total + 1) * MD_WARPS * sizeof(double),
                             S, total, col, i > 0 ? col - Q->ld : (double *)nullptr,
                             i > 1 ? (const double *)(col - 2 * Q->ld) : (const double *)nullptr,
                             (const double *)(a + std::max(i - 1, 0)), (const double *)(b2 + std::max(i - 1, 0)),
                             rows, c.d_partial);
                RAILS_CHECK_LAUNCH();
                rc = rails::finish_reduction(c.d_partial, nb_dot, total + 1, g, reduce, dT, k, coef, b2 + i, nullptr,
                                             alpha_here);
                if (rc) return rc;
            }
            RAILS_LAUNCH(multi_gemv_kernel, nb_gemv, MG_THREADS, (size_t)std::max(total, 1) * sizeof(double),
                         S, total, coef, col + Q->ld, alpha_from_dot ? (const double *)col : (const double *)nullptr,
                         rows, c.d_partial);
            RAILS_CHECK_LAUNCH();
            if (alpha_from_dot)
            {
                rc = rails::finish_reduction(c.d_partial, nb_gemv, 1, a + i, reduce, nullptr, 0, nullptr, nullptr, b2 + i);
                if (rc) return rc;
            }
        }
        // columns 0 .. L-2 were normalised on the way; the last one here
        if (rows > 0)
        {
            double *last = Q->data + (size_t)(qc0 + L - 1) * Q->ld;
            const int nbn = (int)std::max<long long>(1, std::min<long long>((rows + 255) / 256, (long long)c.sm_count * 8));
            RAILS_LAUNCH(normalize_kernel, nbn, 256, 0, last, (const double *)(b2 + L - 1), rows);
            RAILS_CHECK_LAUNCH();
        }
    }
    // one read-back for the whole recurrence
    rc = rails::ensure_small(2 * (size_t)LZ_MAXL + 1);
    if (rc) return rc;
    RAILS_CUDA_TRY(cudaMemcpyAsync(c.h_small, b2, (2 * (size_t)LZ_MAXL + 1) * sizeof(double), cudaMemcpyDeviceToHost, c.stream));
    RAILS_CUDA_TRY(cudaStreamSynchronize(c.stream));
    { int xrc = rails::check_exchange_error("lanczos_run"); if (xrc) return xrc; }
    int done = 0;
    for (int i = 0; i < L; ++i)
    {
        alpha[i] = c.h_small[LZ_MAXL + 1 + i];
        beta[i] = std::sqrt(std::fabs(c.h_small[i + 1]));
        done = i + 1;
        if (beta[i] < 1e-14)
            break;
    }
    *steps = done;
    return RAILS_OK;
}

// Starts the recurrence early: the first projection [X_0 X_1 ..]' w_0 and |w_0|^2
// do not depend on T, so they can run on the GPU while the host solves the
// projected problem (src/LyapunovSolver.hpp:209 -> :215).  The matching
// rails_mv_lanczos_run (same Q, qc0 and blocks) picks the result up.
extern "C" int rails_mv_lanczos_begin(int nblk, const rails_mv_t *X, const int *xc0, const int *xn,
                                      rails_mv_t Q, int qc0)
{
    RAILS_REQUIRE_READY();
    rails::Context &c = ctx();
    c.lz_pending = false;
    if (!Q || qc0 < 0 || qc0 >= Q->cols)
        return rails::fail(RAILS_ERR_ARG, "lanczos_begin: bad start column");
    Segs S;
    int total = 0;
    bool dist = false;
    int rc = collect_segs(nblk, X, xc0, xn, Q->rows, &S, &total, &dist, "lanczos_begin");
    if (rc) return rc;
    if (total > MD_MAXCOLS)
        return RAILS_OK;                 // nothing started; lanczos_run does everything
    const bool reduce = dist && Q->flags == RAILS_MV_DISTRIBUTED && c.comm;
    rc = lz_scratch(total);
    if (rc) return rc;
    const int tot1 = std::max(total, 1) + 1;
    double *b2 = c.d_lz, *g = b2 + 2 * LZ_MAXL + 1;
    (void)tot1;
    const long long rows = Q->rows;
    const long long ntiles = (rows + MD_TILE - 1) / MD_TILE;
    const int nb_dot = (int)std::max<long long>(1, std::min<long long>(ntiles, (long long)c.sm_count * 4));
    rc = rails::ensure_partial((size_t)nb_dot * (total + 1));
    if (rc) return rc;
    rc = lz_attr();
    if (rc) return rc;
    double *col = Q->data + (size_t)qc0 * Q->ld;
    RAILS_LAUNCH(lanczos_dot_kernel, nb_dot, MD_THREADS, (size_t)(total + 1) * MD_WARPS * sizeof(double),
                 S, total, col, (double *)nullptr, (const double *)nullptr, (const double *)b2, (const double *)b2,
                 rows, c.d_partial);
    RAILS_CHECK_LAUNCH();
    rc = rails::finish_reduction(c.d_partial, nb_dot, total + 1, g, reduce, nullptr, 0, nullptr, b2);
    if (rc) return rc;
    c.lz_pending = true;
    c.lz_q = col;
    c.lz_total = total;
    return RAILS_OK;
}

// ---- timing of the step kernel alone (see StepTiming) --------------------------------
extern "C" int rails_lanczos_step_timing(int enable)
{
    g_step_timing.on = enable != 0;
    g_step_timing.used = 0;
    return RAILS_OK;
}

extern "C" int rails_lanczos_step_timing_read(double *ms_total, int *launches, int *shape5)
{
    StepTiming &t = g_step_timing;
    RAILS_CUDA_TRY(cudaStreamSynchronize(rails::ctx().stream));
    double total = 0.0;
    for (size_t i = 0; i + 1 < t.used; i += 2)
    {
        float ms = 0.f;
        RAILS_CUDA_TRY(cudaEventElapsedTime(&ms, t.ev[i], t.ev[i + 1]));
        total += ms;
    }
    if (ms_total) *ms_total = total;
    if (launches) *launches = (int)(t.used / 2);
    if (shape5)
    {
        shape5[0] = t.cols; shape5[1] = t.tile_rows; shape5[2] = t.stages; shape5[3] = t.groups; shape5[4] = t.ctas;
    }
    t.used = 0;
    return RAILS_OK;
}

// Widest [V AV B] (2k + n columns) the fused Lanczos entry points take; the host loop
// falls back to the operation-by-operation sequence beyond it.
extern "C" int rails_mv_lanczos_max_columns(void) { return MD_MAXCOLS; }

// Host-only view of how the one-kernel Lanczos step would be launched for blocks of
// xn[0..nblk) columns (no CUDA call; tests/test_abi.py checks the invariants on CPU):
// shape6 = {tile rows (0: the two-kernel chain runs instead), pipeline stages, column
// groups, dynamic shared memory in bytes, TMA requests per 128-row slab, consumer warps}.
extern "C" int rails_lanczos_step_shape(int nblk, const int *xn, int *shape6)
{
    if (nblk < 1 || nblk > MAX_SEGS || !xn || !shape6)
        return RAILS_ERR_ARG;
    int n[MAX_SEGS], nseg = 0;
    for (int b = 0; b < nblk; ++b)
        if (xn[b] > 0) n[nseg++] = xn[b];
    StepLayout L = ls_layout(n, nseg);
    int tr = 0, st = 0;
    if (L.total > 0)
        ls_shape(L, &tr, &st);
    int nreq = 1;
    for (int b = 0; b < L.nseg; ++b)
        nreq += (L.n[b] + L.bc[b] - 1) / L.bc[b];
    const int groups = L.total > 0 && L.total <= LS_MAX_COLS ? ls_groups(L.total) : 0;
    shape6[0] = tr;
    shape6[1] = st;
    shape6[2] = groups;
    shape6[3] = tr ? (int)ls_smem_bytes(L, tr, st) : 0;
    shape6[4] = nreq;
    shape6[5] = tr ? tr / 32 * groups : 0;
    return RAILS_OK;
}
