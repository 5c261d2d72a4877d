// Tall-skinny inner products  out = X' Y  (a x b)  and the Gram-Schmidt chain.
//
// Replaces StlWrapper::dot (src/StlWrapper.cpp:394-412, dgemm 'T','N') and
// EpetraMV::dot (src/Epetra_MultiVectorWrapper.cpp:418-438).  In the solver
// these are k x 1 products (V'q, AV'q in resid_lanczos, the CGS2 passes) and
// k x e / e x k projection blocks: 0.25 flop/B, i.e. a pure HBM stream of
// 8 m (a + b) bytes.
//
// Kernel shape: the operand with FEW columns (<= BT per launch) is held in
// registers for a tile of 512*RP rows; the operand with MANY columns is
// streamed once, U columns (U*RP 128-bit loads per thread) in flight.  Each
// column's tile contribution is reduced inside the warp with shuffles and
// accumulated in a per-warp shared-memory slice, so there are no atomics and
// the summation order is fixed: row tiles are dealt round-robin to the
// resident CTAs, a second kernel adds the per-CTA partials in CTA order.
#include "internal.h"

#include <algorithm>
#include <vector>

using rails::ctx;

namespace rails
{
int scal_inv_norm_device(double *x, long long ld, long long rows, const double *d);
int cgs2_fused(rails_mv_s *V, int first_col, int against, int vcol, const double *c1_dev, bool dist);
}

namespace
{

constexpr int DOT_THREADS = 256;
constexpr int DOT_WARPS = DOT_THREADS / 32;
// streamed columns handled per blockIdx.y: keeps the per-warp accumulators
// at 64 KB of shared memory per CTA
__host__ __device__ constexpr int dot_achunk(int bt) { return 1024 / bt; }

template<bool FULL>
__device__ __forceinline__ double2 load2(const double *__restrict__ p, long long r, long long rows)
{
    if (FULL)
        return __ldg(reinterpret_cast<const double2 *>(p + r));
    double2 v = make_double2(0.0, 0.0);
    if (r < rows) v.x = __ldg(p + r);
    if (r + 1 < rows) v.y = __ldg(p + r + 1);
    return v;
}

template<int RP, int BT, int U, bool FULL>
__device__ __forceinline__ void dot_tile(const double *__restrict__ X, long long ldx, int ac,
                                         const double *__restrict__ Y, long long ldy,
                                         long long rows, long long base, double *wacc, int lane)
{
    double2 yv[BT][RP];
#pragma unroll
    for (int t = 0; t < BT; ++t)
#pragma unroll
        for (int v = 0; v < RP; ++v)
            yv[t][v] = load2<FULL>(Y + (size_t)t * ldy, base + v * 512, rows);

    for (int j = 0; j < ac; j += U)
    {
        double2 xv[U][RP];
#pragma unroll
        for (int u = 0; u < U; ++u)
        {
            // columns past the chunk re-read the last one; result discarded
            const int jj = (j + u < ac) ? (j + u) : (ac - 1);
#pragma unroll
            for (int v = 0; v < RP; ++v)
                xv[u][v] = load2<FULL>(X + (size_t)jj * ldx, base + v * 512, rows);
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
        {
            double p[BT];
#pragma unroll
            for (int t = 0; t < BT; ++t)
            {
                double s = 0.0;
#pragma unroll
                for (int v = 0; v < RP; ++v)
                {
                    s = fma(xv[u][v].x, yv[t][v].x, s);
                    s = fma(xv[u][v].y, yv[t][v].y, s);
                }
                p[t] = s;
            }
#pragma unroll
            for (int off = 16; off > 0; off >>= 1)
#pragma unroll
                for (int t = 0; t < BT; ++t)
                    p[t] += __shfl_xor_sync(0xffffffffu, p[t], off);
            if (lane == 0 && j + u < ac)
            {
#pragma unroll
                for (int t = 0; t < BT; ++t)
                    wacc[(j + u) * BT + t] += p[t];
            }
        }
    }
}

// partial[blockIdx.x][(j0 + j) * BT + t] = sum over this CTA's row tiles of
// X[:, j0 + j] . Y[:, t]
template<int RP, int BT, int U>
__global__ void __launch_bounds__(DOT_THREADS)
dot_kernel(const double *__restrict__ X, long long ldx, int a,
           const double *__restrict__ Y, long long ldy, long long rows,
           double *__restrict__ partial)
{
    extern __shared__ double sacc[];
    constexpr int ACH = dot_achunk(BT);
    const int j0 = blockIdx.y * ACH;
    const int ac = min(ACH, a - j0);
    const int nacc = ac * BT;
    for (int i = threadIdx.x; i < nacc * DOT_WARPS; i += DOT_THREADS)
        sacc[i] = 0.0;
    __syncthreads();

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    double *wacc = sacc + warp * nacc;
    const double *Xc = X + (size_t)j0 * ldx;
    constexpr long long TILE = 512LL * RP;
    const long long ntiles = (rows + TILE - 1) / TILE;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x)
    {
        const long long base = tile * TILE + 2 * threadIdx.x;
        if ((tile + 1) * TILE <= rows)
            dot_tile<RP, BT, U, true>(Xc, ldx, ac, Y, ldy, rows, base, wacc, lane);
        else
            dot_tile<RP, BT, U, false>(Xc, ldx, ac, Y, ldy, rows, base, wacc, lane);
    }
    __syncthreads();
    double *out = partial + (size_t)blockIdx.x * ((size_t)a * BT) + (size_t)j0 * BT;
    for (int i = threadIdx.x; i < nacc; i += DOT_THREADS)
    {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < DOT_WARPS; ++w)
            s += sacc[w * nacc + i];
        out[i] = s;
    }
}

// One warp per output entry: out = sum_b partial[b][idx], fixed order.
// idx = j * bt + t maps to out[j * sj + (t0 + t) * st].
__global__ void __launch_bounds__(256)
dot_finish_kernel(const double *__restrict__ partial, int nblocks, int n, int bt, int t0,
                  double *__restrict__ out, long long sj, long long st)
{
    const int idx = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (idx >= n) return;
    const int lane = threadIdx.x & 31;
    double s = 0.0;
    for (int b = lane; b < nblocks; b += 32)
        s += partial[(size_t)b * n + idx];
    for (int off = 16; off > 0; off >>= 1)
        s += __shfl_xor_sync(0xffffffffu, s, off);
    if (lane == 0)
    {
        const int j = idx / bt, t = idx - j * bt;
        out[(size_t)j * sj + (size_t)(t0 + t) * st] = s;
    }
}

template<int RP, int BT, int U>
int launch_dot(const double *X, long long ldx, int a, const double *Y, long long ldy,
               long long rows, int t0, double *d_out, long long sj, long long st)
{
    rails::Context &c = ctx();
    constexpr long long TILE = 512LL * RP;
    long long ntiles = (rows + TILE - 1) / TILE;
    if (ntiles < 1) ntiles = 1;
    int nb = (int)std::min<long long>(ntiles, (long long)c.sm_count * 4);
    int rc = rails::ensure_partial((size_t)nb * a * BT);
    if (rc) return rc;
    constexpr int ACH = dot_achunk(BT);
    dim3 grid(nb, (a + ACH - 1) / ACH);
    size_t smem = (size_t)std::min(a, ACH) * BT * DOT_WARPS * sizeof(double);
    static int configured_device = -1;
    if (configured_device != c.device)
    {
        RAILS_CUDA_TRY(cudaFuncSetAttribute(dot_kernel<RP, BT, U>,
                                            cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            ACH * BT * DOT_WARPS * (int)sizeof(double)));
        configured_device = c.device;
    }
    auto kern = dot_kernel<RP, BT, U>;
    RAILS_LAUNCH(kern, grid, DOT_THREADS, smem, X, ldx, a, Y, ldy, rows, c.d_partial);
    RAILS_CHECK_LAUNCH();
    int n = a * BT;
    RAILS_LAUNCH(dot_finish_kernel, (n + 7) / 8, 256, 0, c.d_partial, nb, n, BT, t0, d_out, sj, st);
    RAILS_CHECK_LAUNCH();
    return RAILS_OK;
}

}

namespace rails
{

// d_out (device, column-major a x b, leading dimension ldo) = X' Y over this
// rank's rows, summed over ranks when asked.
int dot_device(const double *X, long long ldx, int a, const double *Y, long long ldy, int b,
               long long rows, double *d_out, int ldo, bool reduce_over_ranks)
{
    Context &c = ctx();
    if (a == 0 || b == 0)
        return RAILS_OK;
    if (rows == 0)
    {
        RAILS_CUDA_TRY(cudaMemset2DAsync(d_out, (size_t)ldo * sizeof(double), 0,
                                         (size_t)a * sizeof(double), b, c.stream));
    }
    else if (fp64_tensor_enabled() && std::min(a, b) > 8 && ((size_t)X % 16 == 0) && ((size_t)Y % 16 == 0) && ldx % 2 == 0 &&
             ldy % 2 == 0)
    {
        // both blocks wide: FP64-arithmetic-bound, tensor-core kernel (gram.cu)
        int rc = gram_device(X, ldx, a, Y, ldy, b, rows, d_out, ldo);
        if (rc) return rc;
    }
    else
    {
        // stream the operand with more columns (S), hold the other (H) in registers
        const bool swap = b > a;
        const double *S = swap ? Y : X, *H = swap ? X : Y;
        const long long lds = swap ? ldy : ldx, ldh = swap ? ldx : ldy;
        const int ns = swap ? b : a, nh = swap ? a : b;
        // out(i, j): streamed column index moves with sj, held with st
        const long long sj = swap ? ldo : 1, st = swap ? 1 : ldo;
        int t = 0;
        while (t < nh)
        {
            const int left = nh - t;
            const double *Ht = H + (size_t)t * ldh;
            int rc;
            if (left >= 8)      { rc = launch_dot<1, 8, 4>(S, lds, ns, Ht, ldh, rows, t, d_out, sj, st); t += 8; }
            else if (left >= 4) { rc = launch_dot<2, 4, 2>(S, lds, ns, Ht, ldh, rows, t, d_out, sj, st); t += 4; }
            else if (left >= 2) { rc = launch_dot<4, 2, 2>(S, lds, ns, Ht, ldh, rows, t, d_out, sj, st); t += 2; }
            else                { rc = launch_dot<4, 1, 4>(S, lds, ns, Ht, ldh, rows, t, d_out, sj, st); t += 1; }
            if (rc) return rc;
        }
    }
    if (reduce_over_ranks && c.comm)
    {
        if (ldo != a)
            return fail(RAILS_ERR_ARG, "dot_device: rank reduction needs a compact result");
        int rc = finish_reduction(d_out, 1, a * b, d_out, true);
        if (rc) return rc;
    }
    return RAILS_OK;
}

}

namespace
{

int check_pair(rails_mv_t X, int xc0, int a, rails_mv_t Y, int yc0, int b, const char *what)
{
    if (!X || !Y)
        return rails::fail(RAILS_ERR_ARG, "%s: null multivector", what);
    if (xc0 < 0 || a < 0 || xc0 + a > X->cols || yc0 < 0 || b < 0 || yc0 + b > Y->cols)
        return rails::fail(RAILS_ERR_ARG, "%s: column range outside capacity", what);
    if (X->rows != Y->rows)
        return rails::fail(RAILS_ERR_ARG, "%s: %lld vs %lld rows", what, X->rows, Y->rows);
    return RAILS_OK;
}

}

extern "C" int rails_mv_dot(rails_mv_t X, int xc0, int a, rails_mv_t Y, int yc0, int b,
                            double *out, int ldo)
{
    RAILS_REQUIRE_READY();
    int rc = check_pair(X, xc0, a, Y, yc0, b, "dot");
    if (rc) return rc;
    if (a == 0 || b == 0) return RAILS_OK;
    if (!out || ldo < a)
        return rails::fail(RAILS_ERR_ARG, "dot: bad result block");
    rails::Context &c = ctx();
    rc = rails::ensure_small((size_t)a * b);
    if (rc) return rc;
    const bool dist = X->flags == RAILS_MV_DISTRIBUTED && Y->flags == RAILS_MV_DISTRIBUTED;
    rc = rails::dot_device(X->data + (size_t)xc0 * X->ld, X->ld, a, Y->data + (size_t)yc0 * Y->ld,
                           Y->ld, b, X->rows, c.d_small, a, dist);
    if (rc) return rc;
    RAILS_CUDA_TRY(cudaMemcpyAsync(c.h_small, c.d_small, (size_t)a * b * sizeof(double),
                                   cudaMemcpyDeviceToHost, c.stream));
    RAILS_CUDA_TRY(cudaStreamSynchronize(c.stream));
    { int xrc = rails::check_exchange_error("dot"); if (xrc) return xrc; }
    for (int j = 0; j < b; ++j)
        for (int i = 0; i < a; ++i)
            out[i + (size_t)j * ldo] = c.h_small[i + (size_t)j * a];
    return RAILS_OK;
}

// Several inner products with ONE host synchronisation: all are enqueued, the
// results of the distributed pairs are all-reduced as one buffer, and a single
// read-back returns everything (the projection update of
// src/LyapunovSolver.hpp:173-195 issues up to six dots in a row).
extern "C" int rails_mv_dot_batch(int count, const rails_mv_t *X, const int *xc0, const int *a,
                                  const rails_mv_t *Y, const int *yc0, const int *b,
                                  double *const *out, const int *ldo)
{
    RAILS_REQUIRE_READY();
    if (count < 0 || (count > 0 && (!X || !Y || !xc0 || !a || !yc0 || !b || !out || !ldo)))
        return rails::fail(RAILS_ERR_ARG, "dot_batch: null argument");
    size_t total = 0;
    for (int i = 0; i < count; ++i)
    {
        int rc = check_pair(X[i], xc0[i], a[i], Y[i], yc0[i], b[i], "dot_batch");
        if (rc) return rc;
        if (a[i] && b[i] && (!out[i] || ldo[i] < a[i]))
            return rails::fail(RAILS_ERR_ARG, "dot_batch: bad result block %d", i);
        total += (size_t)a[i] * b[i];
    }
    if (total == 0) return RAILS_OK;
    rails::Context &c = ctx();
    int rc = rails::ensure_small(total);
    if (rc) return rc;
    // distributed pairs first (one contiguous all-reduce), then the replicated ones
    std::vector<size_t> off(count, 0);
    size_t pos = 0, ndist = 0;
    for (int pass = 0; pass < 2; ++pass)
    {
        for (int i = 0; i < count; ++i)
        {
            const bool dist = X[i]->flags == RAILS_MV_DISTRIBUTED && Y[i]->flags == RAILS_MV_DISTRIBUTED;
            if (dist != (pass == 0) || a[i] == 0 || b[i] == 0) continue;
            off[i] = pos;
            rc = rails::dot_device(X[i]->data + (size_t)xc0[i] * X[i]->ld, X[i]->ld, a[i],
                                   Y[i]->data + (size_t)yc0[i] * Y[i]->ld, Y[i]->ld, b[i], X[i]->rows,
                                   c.d_small + pos, a[i], false);
            if (rc) return rc;
            pos += (size_t)a[i] * b[i];
        }
        if (pass == 0) ndist = pos;
    }
    if (ndist && c.comm)
    {
        rc = rails::finish_reduction(c.d_small, 1, (int)ndist, c.d_small, true);
        if (rc) return rc;
    }
    RAILS_CUDA_TRY(cudaMemcpyAsync(c.h_small, c.d_small, total * sizeof(double), cudaMemcpyDeviceToHost, c.stream));
    RAILS_CUDA_TRY(cudaStreamSynchronize(c.stream));
    { int xrc = rails::check_exchange_error("dot"); if (xrc) return xrc; }
    for (int i = 0; i < count; ++i)
        for (int j = 0; j < b[i]; ++j)
            for (int r = 0; r < a[i]; ++r)
                out[i][r + (size_t)j * ldo[i]] = c.h_small[off[i] + r + (size_t)j * a[i]];
    return RAILS_OK;
}

extern "C" int rails_mv_dot_dev(rails_mv_t X, int xc0, int a, rails_mv_t Y, int yc0, int b,
                                rails_mv_t Out, int oc0)
{
    RAILS_REQUIRE_READY();
    int rc = check_pair(X, xc0, a, Y, yc0, b, "dot_dev");
    if (rc) return rc;
    if (!Out || oc0 < 0 || oc0 + b > Out->cols || Out->rows != a)
        return rails::fail(RAILS_ERR_ARG, "dot_dev: result block must be %d x %d", a, b);
    if (a == 0 || b == 0) return RAILS_OK;
    rails::Context &c = ctx();
    rc = rails::ensure_small((size_t)a * b);
    if (rc) return rc;
    const bool dist = X->flags == RAILS_MV_DISTRIBUTED && Y->flags == RAILS_MV_DISTRIBUTED;
    rc = rails::dot_device(X->data + (size_t)xc0 * X->ld, X->ld, a, Y->data + (size_t)yc0 * Y->ld,
                           Y->ld, b, X->rows, c.d_small, a, dist);
    if (rc) return rc;
    RAILS_CUDA_TRY(cudaMemcpy2DAsync(Out->data + (size_t)oc0 * Out->ld, Out->ld * sizeof(double),
                                     c.d_small, (size_t)a * sizeof(double),
                                     (size_t)a * sizeof(double), b, cudaMemcpyDeviceToDevice, c.stream));
    return RAILS_OK;
}

// CGS2 (StlWrapper::orthogonalize, src/StlWrapper.cpp:305-321):
//   for i in [first, last]:  v = V(:, i);  v /= |v|;
//       twice:  v -= V(:, 0:i-1) * (V(:, 0:i-1)' v);      v /= |v|
// Every scalar (the squared norms, the k x 1 coefficient vectors) stays in
// device memory between kernels; with a communicator the sums are reduced over
// the ranks on the same stream.  Nothing returns to the host.
//
// Several new columns are processed in panels of 8: a panel is first projected,
// twice, against ALL columns before it as one block (two k x 8 products instead
// of 8 x two k x 1: V is streamed twice per panel, not twice per column), then
// its columns are orthogonalised among themselves column by column.  In exact
// arithmetic this is the column-wise recurrence; in floating point it differs
// from it by rounding only (after the two block passes the panel is orthogonal
// to the old columns to working precision).  RAILS_BLOCK_CGS2=0 forces the
// column-by-column sequence.
extern "C" int rails_mv_orthogonalize(rails_mv_t V, int col0, int first, int last)
{
    RAILS_REQUIRE_READY();
    if (!V || col0 < 0 || first < 0 || last < first - 1 || col0 + last + 1 > V->cols)
        return rails::fail(RAILS_ERR_ARG, "orthogonalize: bad column range");
    rails::Context &c = ctx();
    int rc = rails::ensure_small((size_t)last + 2);
    if (rc) return rc;
    const bool dist = V->flags == RAILS_MV_DISTRIBUTED;
    double *base = V->data + (size_t)col0 * V->ld;
    static int block_mode = -1;
    if (block_mode < 0)
    {
        const char *e = getenv("RAILS_BLOCK_CGS2");
        block_mode = (e && e[0] == '0') ? 0 : 1;
    }
    const int PANEL = 8;
    auto normalise = [&](double *v) -> int {
        int r = rails::dot_device(v, V->ld, 1, v, V->ld, 1, V->rows, c.d_small, 1, dist);
        if (r) return r;
        return rails::scal_inv_norm_device(v, V->ld, V->rows, c.d_small);
    };
    int p0 = first;
    while (p0 <= last)
    {
        const int p1 = (block_mode && last > first) ? std::min(last, p0 + PANEL - 1) : p0;
        const int e = p1 - p0 + 1;
        const int old = (e > 1) ? p0 : 0;      // columns handled by the block passes
        double *W = base + (size_t)p0 * V->ld;
        if (e > 1)
        {
            rc = rails::ensure_small((size_t)std::max(p0, 1) * e + 2);
            if (rc) return rc;
            for (int i = 0; i < e; ++i)
            {
                rc = normalise(W + (size_t)i * V->ld);
                if (rc) return rc;
            }
            if (p0 > 0)
                for (int pass = 0; pass < 2; ++pass)
                {
                    rc = rails::dot_device(base, V->ld, p0, W, V->ld, e, V->rows, c.d_small, p0, dist);
                    if (rc) return rc;
                    rc = rails::gemm_device(base, V->ld, p0, c.d_small, p0, e, -1.0, 1.0, W, V->ld, V->rows);
                    if (rc) return rc;
                }
        }
        for (int i = p0; i <= p1; ++i)
        {
            double *v = base + (size_t)i * V->ld;
            if (e == 1)
            {
                rc = normalise(v);
                if (rc) return rc;
            }
            const int against = i - old;        // columns [old, i)
            if (against > 0 && e == 1)
            {
                // one column: first projection, then both updates, the second projection
                // and the norm out of two step-kernel passes (lanczos.cu: cgs2_fused)
                const double *Vo1 = base + (size_t)old * V->ld;
                rc = rails::dot_device(Vo1, V->ld, against, v, V->ld, 1, V->rows, c.d_small, against, dist);
                if (rc) return rc;
                rc = rails::cgs2_fused(V, col0 + old, against, col0 + i, c.d_small, dist);
                if (rc == RAILS_OK) continue;
                if (rc != 1) return rc;
                // (shape outside the step kernel: the plain chain below; its first dot is repeated)
            }
            if (against > 0)
            {
                const double *Vo = base + (size_t)old * V->ld;
                for (int pass = 0; pass < 2; ++pass)
                {
                    rc = rails::dot_device(Vo, V->ld, against, v, V->ld, 1, V->rows, c.d_small, against, dist);
                    if (rc) return rc;
                    rc = rails::gemm_device(Vo, V->ld, against, c.d_small, against, 1, -1.0, 1.0, v, V->ld, V->rows);
                    if (rc) return rc;
                }
            }
            rc = normalise(v);
            if (rc) return rc;
        }
        if (e > 1 && p0 > 0)
        {
            // A column that (nearly) lies in the span of the columns before it in
            // its own panel -- two Ritz vectors that are copies of one another --
            // collapses to rounding noise in the loop above, and normalising noise
            // brings back O(1) components along the OLD columns.  The column-wise
            // recurrence is immune (its second pass runs against all columns); here
            // one more block pass against the old columns restores the same
            // guarantee, followed by a re-normalisation and one intra-panel pass
            // (which only see rounding-size corrections for healthy columns).
            rc = rails::dot_device(base, V->ld, p0, W, V->ld, e, V->rows, c.d_small, p0, dist);
            if (rc) return rc;
            rc = rails::gemm_device(base, V->ld, p0, c.d_small, p0, e, -1.0, 1.0, W, V->ld, V->rows);
            if (rc) return rc;
            for (int i = p0; i <= p1; ++i)
            {
                double *v = base + (size_t)i * V->ld;
                if (i > p0)
                {
                    rc = rails::dot_device(W, V->ld, i - p0, v, V->ld, 1, V->rows, c.d_small, i - p0, dist);
                    if (rc) return rc;
                    rc = rails::gemm_device(W, V->ld, i - p0, c.d_small, i - p0, 1, -1.0, 1.0, v, V->ld, V->rows);
                    if (rc) return rc;
                }
                rc = normalise(v);
                if (rc) return rc;
            }
        }
        p0 = p1 + 1;
    }
    return RAILS_OK;
}

// ---- allocation-free variant -----------------------------------------------------
// rails_mv_dot_dev_noalloc: the streaming X'Y of rails_mv_dot_dev with the per-CTA
// partials in a STATIC device array instead of the grown-on-demand workspace, and
// the result written straight into the caller's device block.  It calls no
// cudaMalloc / cudaFree / allocating routine, does not synchronise and returns
// nothing to the host, so it may be called where allocation is forbidden (CUDA
// graph capture, serving loops).  No counterpart in the
// reference.
//
// The array is part of the library image: every GPU that loads librails_cuda.so
// holds RAILS_NOALLOC_SCRATCH_DOUBLES * 8 bytes (1.16 MiB) until the process exits.
// All calls share it and it keeps the last call's contents.  Nothing in it has to
// start at zero (dot_kernel writes every partial the finish kernel reads), but two
// calls that OVERLAP on the GPU -- issued from different streams or processes'
// threads without ordering -- race on it.  Calls on the library stream are ordered.
namespace
{

constexpr size_t NOALLOC_SCRATCH_DOUBLES = RAILS_NOALLOC_SCRATCH_DOUBLES;
static __device__ __align__(256) double g_noalloc_scratch[NOALLOC_SCRATCH_DOUBLES];

// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address:
// the address a kernel may use comes from cudaGetSymbolAddress, once per device,
// with that device current.
double *g_noalloc_ptr = nullptr;
int g_noalloc_device = -1;

int noalloc_scratch(double **out)
{
    rails::Context &c = ctx();
    if (g_noalloc_device != c.device || !g_noalloc_ptr)
    {
        RAILS_CUDA_TRY(cudaSetDevice(c.device));
        void *p = nullptr;
        RAILS_CUDA_TRY(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        g_noalloc_ptr = static_cast<double *>(p);
        g_noalloc_device = c.device;
    }
    *out = g_noalloc_ptr;
    return RAILS_OK;
}

// CTAs of one launch (same rule as launch_dot) and the scratch it needs
template<int RP>
int noalloc_blocks(long long rows)
{
    constexpr long long TILE = 512LL * RP;
    long long ntiles = std::max<long long>((rows + TILE - 1) / TILE, 1);
    return (int)std::min<long long>(ntiles, (long long)ctx().sm_count * 4);
}

template<int RP, int BT, int U>
int launch_dot_noalloc(const double *X, long long ldx, int a, const double *Y, long long ldy,
                       long long rows, int t0, double *scratch, double *d_out, long long sj, long long st)
{
    rails::Context &c = ctx();
    const int nb = noalloc_blocks<RP>(rows);
    constexpr int ACH = dot_achunk(BT);
    dim3 grid(nb, (a + ACH - 1) / ACH);
    size_t smem = (size_t)std::min(a, ACH) * BT * DOT_WARPS * sizeof(double);
    static int configured_device = -1;
    if (configured_device != c.device)
    {
        RAILS_CUDA_TRY(cudaFuncSetAttribute(dot_kernel<RP, BT, U>,
                                            cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            ACH * BT * DOT_WARPS * (int)sizeof(double)));
        configured_device = c.device;
    }
    auto kern = dot_kernel<RP, BT, U>;
    RAILS_LAUNCH(kern, grid, DOT_THREADS, smem, X, ldx, a, Y, ldy, rows, scratch);
    RAILS_CHECK_LAUNCH();
    const int n = a * BT;
    RAILS_LAUNCH(dot_finish_kernel, (n + 7) / 8, 256, 0, (const double *)scratch, nb, n, BT, t0, d_out, sj, st);
    RAILS_CHECK_LAUNCH();
    return RAILS_OK;
}

// zero fill of an a x b block with leading dimension ld (rows == 0)
__global__ void noalloc_zero_kernel(double *out, int a, int b, long long ld)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < a * b; i += gridDim.x * blockDim.x)
        out[(i % a) + (size_t)(i / a) * ld] = 0.0;
}

}

extern "C" int rails_mv_dot_dev_noalloc(rails_mv_t X, int xc0, int a, rails_mv_t Y, int yc0, int b,
                                        rails_mv_t Out, int oc0)
{
    RAILS_REQUIRE_READY();
    int rc = check_pair(X, xc0, a, Y, yc0, b, "dot_dev_noalloc");
    if (rc) return rc;
    if (!Out || oc0 < 0 || oc0 + b > Out->cols || Out->rows != a)
        return rails::fail(RAILS_ERR_ARG, "dot_dev_noalloc: result block must be %d x %d", a, b);
    if (a == 0 || b == 0) return RAILS_OK;
    rails::Context &c = ctx();
    if (c.comm && c.nranks > 1 && X->flags == RAILS_MV_DISTRIBUTED && Y->flags == RAILS_MV_DISTRIBUTED)
        return rails::fail(RAILS_ERR_ARG, "dot_dev_noalloc: row-partitioned operands need the sum over ranks; "
                           "use rails_mv_dot_dev");
    const long long rows = X->rows;
    const double *Xd = X->data + (size_t)xc0 * X->ld, *Yd = Y->data + (size_t)yc0 * Y->ld;
    double *out = Out->data + (size_t)oc0 * Out->ld;
    const long long ldo = Out->ld;
    if (rows == 0)
    {
        RAILS_LAUNCH(noalloc_zero_kernel, 1, 256, 0, out, a, b, ldo);
        RAILS_CHECK_LAUNCH();
        return RAILS_OK;
    }
    // stream the operand with more columns, hold the other in registers (as dot_device)
    const bool swap = b > a;
    const double *S = swap ? Yd : Xd, *H = swap ? Xd : Yd;
    const long long lds = swap ? Y->ld : X->ld, ldh = swap ? X->ld : Y->ld;
    const int ns = swap ? b : a, nh = swap ? a : b;
    const long long sj = swap ? ldo : 1, st = swap ? 1 : ldo;
    // the largest launch decides: partials = CTAs x streamed columns x held columns per launch
    size_t need = 0;
    for (int t = 0; t < nh;)
    {
        const int left = nh - t;
        const int bt = left >= 8 ? 8 : left >= 4 ? 4 : left >= 2 ? 2 : 1;
        const int nb = bt == 8 ? noalloc_blocks<1>(rows) : bt == 4 ? noalloc_blocks<2>(rows) : noalloc_blocks<4>(rows);
        need = std::max(need, (size_t)nb * (size_t)ns * (size_t)bt);
        t += bt;
    }
    if (need > NOALLOC_SCRATCH_DOUBLES)
        return rails::fail(RAILS_ERR_ARG, "dot_dev_noalloc: a %d x %d product over %lld rows needs %zu doubles of "
                           "scratch, the static array holds %zu (RAILS_NOALLOC_SCRATCH_DOUBLES); use rails_mv_dot_dev",
                           a, b, rows, need, NOALLOC_SCRATCH_DOUBLES);
    double *scratch = nullptr;
    rc = noalloc_scratch(&scratch);
    if (rc) return rc;
    int t = 0;
    while (t < nh)
    {
        const int left = nh - t;
        const double *Ht = H + (size_t)t * ldh;
        if (left >= 8)      { rc = launch_dot_noalloc<1, 8, 4>(S, lds, ns, Ht, ldh, rows, t, scratch, out, sj, st); t += 8; }
        else if (left >= 4) { rc = launch_dot_noalloc<2, 4, 2>(S, lds, ns, Ht, ldh, rows, t, scratch, out, sj, st); t += 4; }
        else if (left >= 2) { rc = launch_dot_noalloc<4, 2, 2>(S, lds, ns, Ht, ldh, rows, t, scratch, out, sj, st); t += 2; }
        else                { rc = launch_dot_noalloc<4, 1, 4>(S, lds, ns, Ht, ldh, rows, t, scratch, out, sj, st); t += 1; }
        if (rc) return rc;
    }
    return RAILS_OK;
}
