// Block inner products  out = X' Y  (a x b) with BOTH a and b large.
//
// In the solver these are the projection of a warm-start basis V'(AV) (k x k,
// src/LyapunovSolver.hpp:173-195 on the first iteration of a continuation run,
// test/LyapunovSolver_test.cpp:312-352), the k x e / e x k projection blocks with
// e = 16 expansion vectors, and the Gram matrices of the tall-skinny QR and of
// the solution post-processing.  Unlike the k x 1 products of the Lanczos
// recurrence (dot.cu: a pure HBM stream) they carry 2 a b / (8 (a + b)) flop per
// byte -- 8 flop/B at a = b = 64 -- and are bound by FP64 arithmetic.
//
// The streaming kernel of dot.cu holds 8 columns of one operand in registers and
// reduces every product across the warp with shuffles: at 64 x 64 it re-reads X
// eight times and is shuffle-bound (49 ms at m = 16.8 M, 7 % of the FP64 peak).
// Here the product runs on the FP64 tensor cores (DMMA.8x8x4, mma.sync
// m8n8k4.f64): the reduction dimension of the MMA is the ROW index, so the
// fragments are read straight from the column-major operands -- lane (g, t) of a
// warp loads rows r0 + 4t .. r0 + 4t + 3 of column c0 + g as two 128-bit words,
// for A (from X) and B (from Y) alike -- no shared memory, no shuffles, each
// operand read once from HBM (the warps of a CTA share rows through L1).
// A warp owns a 32 x 32 block of the result (16 accumulator fragments), a CTA of
// four warps 64 x 64, 128 x 32 or 32 x 128; CTAs split the rows in contiguous
// chunks and leave per-CTA partial results that a second kernel adds in CTA
// order (no atomics: the summation order is fixed, results are reproducible).
#include "internal.h"

#include <algorithm>
#include <cstdlib>

using rails::ctx;

namespace
{

constexpr int GRAM_THREADS = 128;

__device__ __forceinline__ void dmma(double &c0, double &c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// rows r .. r + 3 of one column; FULL: all four exist and r is a multiple of 2
template<bool FULL>
__device__ __forceinline__ void load4(const double *__restrict__ col, long long r, long long rows,
                                      double (&v)[4])
{
    if (FULL)
    {
        const double2 lo = __ldg(reinterpret_cast<const double2 *>(col + r));
        const double2 hi = __ldg(reinterpret_cast<const double2 *>(col + r + 2));
        v[0] = lo.x; v[1] = lo.y; v[2] = hi.x; v[3] = hi.y;
    }
    else
    {
#pragma unroll
        for (int s = 0; s < 4; ++s)
            v[s] = (r + s < rows) ? __ldg(col + r + s) : 0.0;
    }
}

// WA x WB warps, each with a 32 x 32 block: CTA tile (32 WA) x (32 WB)
template<int WA, int WB>
__global__ void __launch_bounds__(GRAM_THREADS)
gram_dmma_kernel(const double *__restrict__ X, long long ldx, int a,
                 const double *__restrict__ Y, long long ldy, int b, long long rows,
                 long long rows_per_cta, int tiles_b, double *__restrict__ partial)
{
    static_assert(WA * WB * 32 == GRAM_THREADS, "four warps");
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int tile_a = blockIdx.y / tiles_b, tile_b = blockIdx.y - tile_a * tiles_b;
    const int a0 = tile_a * 32 * WA + (warp / WB) * 32;
    const int b0 = tile_b * 32 * WB + (warp % WB) * 32;

    // this lane's four X columns and four Y columns (clamped: results of the
    // columns past the end are never stored)
    const double *xc[4], *yc[4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
    {
        xc[i] = X + (size_t)min(a0 + 8 * i + g, a - 1) * ldx;
        yc[i] = Y + (size_t)min(b0 + 8 * i + g, b - 1) * ldy;
    }

    double acc[4][4][2];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
            acc[i][j][0] = acc[i][j][1] = 0.0;

    const long long begin = (long long)blockIdx.x * rows_per_cta;
    const long long end = min(rows, begin + rows_per_cta);
    const bool active = a0 < a && b0 < b;
    if (active)
    {
        for (long long r0 = begin; r0 < end; r0 += 16)
        {
            double xv[4][4], yv[4][4];
            const long long r = r0 + 4 * t;
            if (r0 + 16 <= end)
            {
#pragma unroll
                for (int i = 0; i < 4; ++i)
                {
                    load4<true>(xc[i], r, rows, xv[i]);
                    load4<true>(yc[i], r, rows, yv[i]);
                }
            }
            else
            {
#pragma unroll
                for (int i = 0; i < 4; ++i)
                {
                    load4<false>(xc[i], r, end, xv[i]);
                    load4<false>(yc[i], r, end, yv[i]);
                }
            }
#pragma unroll
            for (int s = 0; s < 4; ++s)
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        dmma(acc[i][j][0], acc[i][j][1], xv[i][s], yv[j][s]);
        }
    }

    // fragment (i, j): lane holds rows a0 + 8 i + g, columns b0 + 8 j + 2 t, + 1
    double *out = partial + (size_t)blockIdx.x * ((size_t)a * b);
    if (active)
    {
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j)
            {
                const int ia = a0 + 8 * i + g;
                const int jb = b0 + 8 * j + 2 * t;
                if (ia < a && jb < b)
                    out[ia + (size_t)jb * a] = acc[i][j][0];
                if (ia < a && jb + 1 < b)
                    out[ia + (size_t)(jb + 1) * a] = acc[i][j][1];
            }
    }
}

// out[i + ldo j] = sum over CTAs of partial[cta][i + a j], one warp per entry, fixed order
__global__ void __launch_bounds__(256)
gram_finish_kernel(const double *__restrict__ partial, int nblocks, int a, int b,
                   double *__restrict__ out, int ldo)
{
    const int idx = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (idx >= a * b) return;
    const int lane = threadIdx.x & 31;
    double s = 0.0;
    for (int c = lane; c < nblocks; c += 32)
        s += partial[(size_t)c * ((size_t)a * b) + idx];
    for (int off = 16; off > 0; off >>= 1)
        s += __shfl_xor_sync(0xffffffffu, s, off);
    if (lane == 0)
    {
        const int j = idx / a, i = idx - j * a;
        out[i + (size_t)j * ldo] = s;
    }
}

// ---- tall-skinny times small on the tensor cores:  Out = alpha V C + beta Out -------
// V m x k, C k x c with c > 8: the restart compression V <- V X, AV <- AV X
// (src/LyapunovSolver.hpp:265-297) and the block updates of the tall-skinny QR.
// The streaming kernel of gemm.cu re-reads V once per 8 result columns; here a
// warp owns 32 rows x (8 NB) result columns and reads its rows of V ONCE: the MMA
// row index is the row of V, interleaved so that lane (g, t) loads rows
// r0 + 4g .. r0 + 4g + 3 of column k0 + t as two 128-bit words (the A fragments of
// four MMA row blocks) and stores four consecutive rows of a result column.  The
// k x (8 NB) block of C sits in shared memory, zero-padded to a multiple of 4 rows.
constexpr int GEMM_RING = 8;          // k-steps (4 columns x 32 rows = 1 KB each) a warp keeps in flight

template<int NB>
__global__ void __launch_bounds__(GRAM_THREADS, NB <= 4 ? 4 : 2)
gemm_dmma_kernel(const double *__restrict__ V, long long ldv, int k, const double *__restrict__ C,
                 int ldc, int c, double alpha, double beta, double *__restrict__ Out, long long ldo,
                 long long rows, int use_ring)
{
    extern __shared__ double Cs[];
    constexpr int CSLD = NB * 8 + 1;
    const int kpad = (k + 3) & ~3;
    const int n0 = blockIdx.y * NB * 8;
    for (int idx = threadIdx.x; idx < kpad * NB * 8; idx += GRAM_THREADS)
    {
        const int kk = idx % kpad, n = idx / kpad;
        Cs[kk * CSLD + n] = (kk < k && n0 + n < c) ? C[kk + (size_t)(n0 + n) * ldc] : 0.0;
    }
    __syncthreads();

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, t = lane & 3;
    double2 *ring = reinterpret_cast<double2 *>(Cs + (size_t)kpad * CSLD);     // kpad * CSLD * 8 is a multiple of 32
    const long long nblocks = (rows + 31) / 32;
    for (long long rb = (long long)blockIdx.x * 4 + warp; rb < nblocks; rb += (long long)gridDim.x * 4)
    {
        const long long r = rb * 32 + 4 * g;
        const bool full = rb * 32 + 32 <= rows;
        double acc[4][NB][2];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < NB; ++j)
                acc[i][j][0] = acc[i][j][1] = 0.0;
        // software pipeline: the rows of the columns two steps ahead are loaded before
        // the MMAs of this step are issued
        auto fetch = [&](int k0, double (&dst)[4]) {
            const double *vc = V + (size_t)min(k0 + t, k - 1) * ldv;
            if (full)
                load4<true>(vc, r, rows, dst);
            else
                load4<false>(vc, r, rows, dst);
        };
        auto mma_step = [&](int k0, const double (&a)[4]) {
            const double *cs = Cs + (k0 + t) * CSLD + g;
#pragma unroll
            for (int j = 0; j < NB; ++j)
            {
                const double bv = cs[8 * j];
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    dmma(acc[i][j][0], acc[i][j][1], a[i], bv);
            }
        };
        if (full && use_ring)
        {
            // Round 2 (RAILS_GEMM_DMMA_RING=0 selects the register pipeline below; same
            // box, 64 x 32 at 16.8 M rows: 3.22 ms with the ring, 3.76 ms without,
            // profiles/r02_gemm_dmma.md): the rows of V travel through a per-lane ring in shared memory
            // filled by 16-byte asynchronous copies (cp.async.ca), GEMM_RING k-steps deep:
            // the loads of a warp no longer hold registers while in flight (round 1:
            // two steps ahead in registers, 59 % of the DMMA pipe, long-scoreboard
            // stalls), each lane reads back exactly the two words it copied, so the
            // layout is conflict-free and no barrier is needed.
            const int steps = kpad / 4;
            double2 *rw = ring + (size_t)warp * GEMM_RING * 64;
            auto issue = [&](int st) {
                if (st < steps)
                {
                    const double *src = V + (size_t)min(4 * st + t, k - 1) * ldv + r;
                    const int slot = st % GEMM_RING;
                    const unsigned d0 = (unsigned)__cvta_generic_to_shared(rw + (slot * 2) * 32 + lane);
                    const unsigned d1 = (unsigned)__cvta_generic_to_shared(rw + (slot * 2 + 1) * 32 + lane);
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(d0), "l"(src) : "memory");
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(d1), "l"(src + 2) : "memory");
                }
                asm volatile("cp.async.commit_group;" ::: "memory");
            };
            for (int st = 0; st < GEMM_RING - 1; ++st)
                issue(st);
            for (int st = 0; st < steps; ++st)
            {
                issue(st + GEMM_RING - 1);
                asm volatile("cp.async.wait_group %0;" ::"n"(GEMM_RING - 1) : "memory");
                const int slot = st % GEMM_RING;
                const double2 lo = rw[(slot * 2) * 32 + lane], hi = rw[(slot * 2 + 1) * 32 + lane];
                const double a[4] = {lo.x, lo.y, hi.x, hi.y};
                mma_step(4 * st, a);
            }
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        else
        {
            double a0[4], a1[4], a2[4];
            fetch(0, a0);
            fetch(min(4, kpad - 4), a1);
            for (int k0 = 0; k0 < kpad; k0 += 4)
            {
                fetch(min(k0 + 8, kpad - 4), a2);
                mma_step(k0, a0);
#pragma unroll
                for (int i = 0; i < 4; ++i)
                {
                    a0[i] = a1[i];
                    a1[i] = a2[i];
                }
            }
        }
        // fragment (i, j): row r + i, columns n0 + 8 j + 2 t, + 1
#pragma unroll
        for (int j = 0; j < NB; ++j)
#pragma unroll
            for (int e = 0; e < 2; ++e)
            {
                const int col = n0 + 8 * j + 2 * t + e;
                if (col >= c)
                    continue;
                double *oc = Out + (size_t)col * ldo + r;
                double o[4];
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    o[i] = alpha * acc[i][j][e];
                if (full)
                {
                    if (beta != 0.0)
                    {
                        const double2 lo = *reinterpret_cast<const double2 *>(oc);
                        const double2 hi = *reinterpret_cast<const double2 *>(oc + 2);
                        o[0] = fma(beta, lo.x, o[0]); o[1] = fma(beta, lo.y, o[1]);
                        o[2] = fma(beta, hi.x, o[2]); o[3] = fma(beta, hi.y, o[3]);
                    }
                    *reinterpret_cast<double2 *>(oc) = make_double2(o[0], o[1]);
                    *reinterpret_cast<double2 *>(oc + 2) = make_double2(o[2], o[3]);
                }
                else
                {
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        if (r + i < rows)
                            oc[i] = beta != 0.0 ? fma(beta, oc[i], o[i]) : o[i];
                }
            }
    }
}

template<int NB>
int launch_gemm_dmma(const double *V, long long ldv, int k, const double *dC, int ldc, int c,
                     double alpha, double beta, double *Out, long long ldo, long long rows)
{
    rails::Context &cx = ctx();
    const int kpad = (k + 3) & ~3;
    const size_t smem = (size_t)kpad * (NB * 8 + 1) * sizeof(double) +
                        (size_t)(GRAM_THREADS / 32) * GEMM_RING * 64 * sizeof(double2);
    static size_t configured = 0;
    static int configured_device = -1;          // the attribute is per device
    if (configured_device != cx.device)
    {
        configured = 0;
        configured_device = cx.device;
    }
    if (smem > configured)
    {
        RAILS_CUDA_TRY(cudaFuncSetAttribute(gemm_dmma_kernel<NB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)std::max<size_t>(smem, 48 * 1024)));
        configured = smem;
    }
    const long long nblocks = (rows + 31) / 32;
    dim3 grid((unsigned)std::max<long long>(1, std::min<long long>((nblocks + 3) / 4, (long long)cx.sm_count * 6)),
              (unsigned)((c + NB * 8 - 1) / (NB * 8)));
    auto kern = gemm_dmma_kernel<NB>;
    static const int use_ring = std::getenv("RAILS_GEMM_DMMA_RING") ? std::atoi(std::getenv("RAILS_GEMM_DMMA_RING")) : 1;
    RAILS_LAUNCH(kern, grid, GRAM_THREADS, smem, V, ldv, k, dC, ldc, c, alpha, beta, Out, ldo, rows, use_ring);
    RAILS_CHECK_LAUNCH();
    return RAILS_OK;
}

// ---- FP64 peak probes: the denominator of the FLOP-bound rooflines -----------------
__global__ void __launch_bounds__(256)
fp64_fma_probe_kernel(double *out, int iters, double seed)
{
    double x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
        x[i] = seed + i + threadIdx.x;
    const double m = 1.0 - 1e-9, c = 1e-9;
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int i = 0; i < 8; ++i)
            x[i] = fma(x[i], m, c);
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 8; ++i)
        s += x[i];
    if (s == 12345.678)                     // never true: keeps the chain alive
        out[blockIdx.x * 256 + threadIdx.x] = s;
}

__global__ void __launch_bounds__(256)
fp64_mma_probe_kernel(double *out, int iters, double seed)
{
    double acc[8][2];
#pragma unroll
    for (int i = 0; i < 8; ++i)
        acc[i][0] = acc[i][1] = 0.0;
    const double av = seed * 1e-9 + (threadIdx.x & 7) * 1e-12, bv = 1e-3 + (threadIdx.x >> 3) * 1e-9;
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int i = 0; i < 8; ++i)
            dmma(acc[i][0], acc[i][1], av, bv);
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 8; ++i)
        s += acc[i][0] + acc[i][1];
    if (s == 12345.678)
        out[blockIdx.x * 256 + threadIdx.x] = s;
}

template<int WA, int WB>
int launch_gram(const double *X, long long ldx, int a, const double *Y, long long ldy, int b,
                long long rows, double *d_out, int ldo)
{
    rails::Context &c = ctx();
    const int tiles_a = (a + 32 * WA - 1) / (32 * WA), tiles_b = (b + 32 * WB - 1) / (32 * WB);
    const int tiles = tiles_a * tiles_b;
    // enough CTAs to fill the machine three times over, at least 4096 rows each
    long long want = ((long long)c.sm_count * 3 + tiles - 1) / tiles;
    want = std::max<long long>(1, std::min<long long>(want, (rows + 4095) / 4096));
    long long rows_per_cta = (rows + want - 1) / want;
    rows_per_cta = ((rows_per_cta + 15) / 16) * 16;
    const int nb = (int)((rows + rows_per_cta - 1) / rows_per_cta);
    int rc = rails::ensure_partial((size_t)nb * a * b);
    if (rc) return rc;
    dim3 grid(nb, tiles);
    auto kern = gram_dmma_kernel<WA, WB>;
    RAILS_LAUNCH(kern, grid, GRAM_THREADS, 0, X, ldx, a, Y, ldy, b, rows, rows_per_cta, tiles_b,
                 c.d_partial);
    RAILS_CHECK_LAUNCH();
    RAILS_LAUNCH(gram_finish_kernel, (a * b + 7) / 8, 256, 0, c.d_partial, nb, a, b, d_out, ldo);
    RAILS_CHECK_LAUNCH();
    return RAILS_OK;
}

}

namespace rails
{

// RAILS_FP64_TENSOR=0 in the environment keeps every product on the streaming
// kernels of dot.cu / gemm.cu (A/B timing of the two paths)
bool fp64_tensor_enabled()
{
    static const bool on = [] {
        const char *e = std::getenv("RAILS_FP64_TENSOR");
        return !(e && e[0] == '0');
    }();
    return on;
}

// d_out (device, column-major a x b, leading dimension ldo) = X' Y over `rows`
// rows; both operands 16-byte aligned with even leading dimensions (every
// multivector is: ld is a multiple of 16 doubles).
int gram_device(const double *X, long long ldx, int a, const double *Y, long long ldy, int b,
                long long rows, double *d_out, int ldo)
{
    if (b <= 32 && a > 32)
        return launch_gram<4, 1>(X, ldx, a, Y, ldy, b, rows, d_out, ldo);
    if (a <= 32 && b > 32)
        return launch_gram<1, 4>(X, ldx, a, Y, ldy, b, rows, d_out, ldo);
    return launch_gram<2, 2>(X, ldx, a, Y, ldy, b, rows, d_out, ldo);
}

}

namespace rails
{

// Out = alpha V C + beta Out for c > 8 result columns; false when the shape does
// not fit this kernel (the caller falls back to the streaming one)
bool gemm_dmma_fits(int k, int c)
{
    const int nb = c > 32 ? 8 : 4;
    return fp64_tensor_enabled() && c > 8 && k >= 4 &&
           (size_t)((k + 3) & ~3) * (nb * 8 + 1) * sizeof(double) + (size_t)(GRAM_THREADS / 32) * GEMM_RING * 1024 <= 200 * 1024;
}

int gemm_dmma_device(const double *V, long long ldv, int k, const double *dC, int ldc, int c,
                     double alpha, double beta, double *Out, long long ldo, long long rows)
{
    if (c > 32)
        return launch_gemm_dmma<8>(V, ldv, k, dC, ldc, c, alpha, beta, Out, ldo, rows);
    return launch_gemm_dmma<4>(V, ldv, k, dC, ldc, c, alpha, beta, Out, ldo, rows);
}

}

// Sustained FP64 throughput of this GPU, TFLOP/s: kind 0 = DFMA (vector pipe),
// 1 = DMMA.8x8x4 (tensor cores).  A few milliseconds of a register-only kernel on
// every SM, timed with CUDA events on the library stream.
extern "C" int rails_cuda_fp64_peak(int kind, double *tflops)
{
    RAILS_REQUIRE_READY();
    if (!tflops || kind < 0 || kind > 1)
        return rails::fail(RAILS_ERR_ARG, "rails_cuda_fp64_peak(%d)", kind);
    rails::Context &c = ctx();
    const int blocks = c.sm_count * 8, iters = 4096;
    double *scratch = nullptr;
    RAILS_CUDA_TRY(cudaMalloc((void **)&scratch, (size_t)blocks * 256 * sizeof(double)));
    cudaEvent_t e0, e1;
    RAILS_CUDA_TRY(cudaEventCreate(&e0));
    RAILS_CUDA_TRY(cudaEventCreate(&e1));
    float best = 0.0f;
    for (int rep = 0; rep < 4; ++rep)
    {
        RAILS_CUDA_TRY(cudaEventRecord(e0, c.stream));
        if (kind == 0)
            RAILS_LAUNCH(fp64_fma_probe_kernel, blocks, 256, 0, scratch, iters, 1.0 + rep);
        else
            RAILS_LAUNCH(fp64_mma_probe_kernel, blocks, 256, 0, scratch, iters, 1.0 + rep);
        RAILS_CUDA_TRY(cudaEventRecord(e1, c.stream));
        RAILS_CUDA_TRY(cudaEventSynchronize(e1));
        float ms = 0.0f;
        RAILS_CUDA_TRY(cudaEventElapsedTime(&ms, e0, e1));
        if (rep > 0 && (best == 0.0f || ms < best))
            best = ms;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(scratch);
    RAILS_CHECK_LAUNCH();
    // flops: DFMA: threads * 8 chains * iters * 2;  DMMA: warps * 8 fragments * iters * (8*8*4*2)
    const double threads = (double)blocks * 256.0;
    const double flops = kind == 0 ? threads * 8.0 * iters * 2.0 : threads / 32.0 * 8.0 * iters * 512.0;
    *tflops = flops / ((double)best * 1e-3) / 1e12;
    return RAILS_OK;
}
