// Small reductions over the ranks without a library collective.
//
// Every reduction on the hot path ends in a handful of doubles (k x e inner
// products, the scalars of a Lanczos step) that all ranks need.  With NCCL
// that is: a second-stage kernel, an all-reduce kernel (launch + protocol
// latency, ~20 us at 8 ranks), and whatever consumes the result.  Here ONE
// single-CTA kernel does all of it over NVLink/NVSwitch peer memory:
//
//   1. adds the per-CTA partials of the producing kernel in a fixed order;
//   2. stores the n local sums into slot (seq mod 4) of EVERY rank's mailbox
//      (plain 8-byte stores to peer-mapped memory), fences, then releases a
//      flag = seq in each mailbox;
//   3. waits until the P flags of its own mailbox show seq, adds the P
//      contributions in rank order (the same order on every rank, so all ranks
//      hold identical bits -- the replicated k x k host objects stay replicated);
//   4. optionally continues with the k x k products T * g of the Lanczos step.
//
// Slots: a rank can run at most one exchange ahead of the slowest peer (its
// next-but-one store needs the peer's contribution to the next exchange, which
// the peer issues only after it has consumed the current one), so two slots
// would do; four are used.  Flags are monotonic sequence numbers, never reset.
// A wait that exceeds RAILS_MAILBOX_TIMEOUT_S (default 30 s) makes the result NaN
// and raises a host-visible flag instead of hanging the GPU or summing stale slots.
#include "internal.h"

#include <cstring>

using rails::ctx;
using rails::Mailbox;
using rails::MB_MAX_RANKS;
using rails::MB_MAXN;
using rails::MB_SLOTS;

namespace
{

constexpr int FX_THREADS = 1024;

__device__ __forceinline__ void st_release_sys(unsigned long long *p, unsigned long long v)
{
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

__global__ void __launch_bounds__(FX_THREADS)
finish_exchange_kernel(const double *partial, int nblocks, int n, double *out, Mailbox mb,
                       unsigned long long seq, volatile int *error, long long timeout_cycles,
                       const double *__restrict__ T, int k,
                       double *__restrict__ coef, double *norm_out, const double *scale_sq,
                       double *alpha_out)
{
    __shared__ double sm[FX_THREADS];
    extern __shared__ double res[];          // n doubles: the finished result (+ n: coef, with alpha_out)
    const int t = threadIdx.x;

    // ---- 1. partials -> local sums.  Fixed order: `groups` (a power of two)
    // interleaved subsets of the blocks, each summed sequentially, then a tree
    // over the groups.
    int groups = 1;
    if (n < FX_THREADS)
    {
        const int cap = min(FX_THREADS / n, nblocks);
        while (groups * 2 <= cap) groups *= 2;
    }
    if (n < FX_THREADS)
    {
        const int i = t % n, grp = t / n;
        if (grp < groups)
        {
            double s = 0.0;
#pragma unroll 8
            for (int b = grp; b < nblocks; b += groups)
                s += partial[(size_t)b * n + i];
            sm[grp * n + i] = s;
        }
        __syncthreads();
        for (int h = groups / 2; h > 0; h >>= 1)
        {
            if (grp < h)
                sm[grp * n + i] += sm[(grp + h) * n + i];
            __syncthreads();
        }
        if (t < n)
            res[t] = sm[t];
    }
    else
    {
        for (int i = t; i < n; i += FX_THREADS)
        {
            double s = 0.0;
#pragma unroll 8
            for (int b = 0; b < nblocks; ++b)
                s += partial[(size_t)b * n + i];
            res[i] = s;
        }
    }
    __syncthreads();

    // ---- 2./3. exchange over peer memory
    if (mb.nranks > 1)
    {
        const int slot = (int)(seq % MB_SLOTS);
        const size_t mine = ((size_t)slot * MB_MAX_RANKS + mb.rank) * MB_MAXN;
        for (int p = 0; p < mb.nranks; ++p)
        {
            double *dst = mb.data[p] + mine;
            for (int i = t; i < n; i += FX_THREADS)
                dst[i] = res[i];
        }
        __threadfence_system();
        __syncthreads();
        if (t < mb.nranks)
            st_release_sys(mb.flags[t] + slot * MB_MAX_RANKS + mb.rank, seq);
        __shared__ int timed_out;
        if (t == 0) timed_out = 0;
        __syncthreads();
        if (t < mb.nranks)
        {
            const unsigned long long *f = mb.flags[mb.rank] + slot * MB_MAX_RANKS + t;
            const long long t0 = clock64();
            while (ld_acquire_sys(f) < seq)
            {
                if (clock64() - t0 > timeout_cycles)      // a peer died or is far behind
                {
                    timed_out = 1;
                    break;
                }
            }
        }
        __syncthreads();
        const double *src = mb.data[mb.rank] + (size_t)slot * MB_MAX_RANKS * MB_MAXN;
        if (timed_out)
        {
            // never hand out a sum of stale slots: the result is NaN on this rank and
            // the host sees the flag at its next read-back (check_exchange_error)
            if (t == 0)
            {
                *error = 1;
                __threadfence_system();
            }
            for (int i = t; i < n; i += FX_THREADS)
                res[i] = __longlong_as_double(0x7ff8000000000000LL);
        }
        else
            for (int i = t; i < n; i += FX_THREADS)
            {
                double s = 0.0;
                for (int r = 0; r < mb.nranks; ++r)
                    s += ((const volatile double *)src)[(size_t)r * MB_MAXN + i];
                res[i] = s;
            }
        __syncthreads();
    }
    // optional scaling by a squared norm held on the device
    if (norm_out)
    {
        const double sq = res[n - 1];
        __syncthreads();
        if (t == 0) *norm_out = sq;
        n -= 1;
        const double sc = 1.0 / sqrt(fabs(sq));
        for (int i = t; i < n; i += FX_THREADS)
            res[i] *= sc;
        __syncthreads();
    }
    else if (scale_sq)
    {
        const double sc = 1.0 / sqrt(fabs(*scale_sq));
        for (int i = t; i < n; i += FX_THREADS)
            res[i] *= sc;
        __syncthreads();
    }
    for (int i = t; i < n; i += FX_THREADS)
        out[i] = res[i];

    // ---- 4. coef = [T g_1 ; T g_0 ; g_rest]   (src/LyapunovSolver.hpp:395,401)
    if (coef)
    {
        double *cf = res + n + (norm_out ? 1 : 0);       // second half of the dynamic block
        const int warp = t >> 5, lane = t & 31;
        for (int row = warp; row < 2 * k; row += FX_THREADS / 32)
        {
            const int j = row < k ? row : row - k;
            const double *gv = row < k ? res + k : res;
            double s = 0.0;
            for (int l = lane; l < k; l += 32)
                s = fma(T[j + (size_t)l * k], gv[l], s);
            for (int o = 16; o > 0; o >>= 1)
                s += __shfl_xor_sync(0xffffffffu, s, o);
            if (lane == 0)
            {
                coef[row] = s;
                if (alpha_out) cf[row] = s;
            }
        }
        for (int i = 2 * k + t; i < n; i += FX_THREADS)
        {
            coef[i] = res[i];
            if (alpha_out) cf[i] = res[i];
        }
        // ---- 5. alpha = q'R q = sum_b (X_b'q) . coef_b = g . coef (:406: `next.dot(q)`
        // with next = sum_b X_b coef_b), from what this kernel already holds: the
        // step needs no second reduction over the rows and no second exchange
        if (alpha_out)
        {
            __syncthreads();
            double s = 0.0;
            for (int i = t; i < n; i += FX_THREADS)
                s = fma(res[i], cf[i], s);
            sm[t] = s;
            __syncthreads();
            for (int h = FX_THREADS / 2; h > 0; h >>= 1)
            {
                if (t < h) sm[t] += sm[t + h];
                __syncthreads();
            }
            if (t == 0) *alpha_out = sm[0];
        }
    }
}

__global__ void __launch_bounds__(256)
alpha_plain_kernel(const double *__restrict__ g, const double *__restrict__ coef, int n, double *alpha_out)
{
    __shared__ double red[256];
    double s = 0.0;
    for (int i = threadIdx.x; i < n; i += 256)
        s = fma(g[i], coef[i], s);
    red[threadIdx.x] = s;
    __syncthreads();
    for (int h = 128; h > 0; h >>= 1)
    {
        if (threadIdx.x < h) red[threadIdx.x] += red[threadIdx.x + h];
        __syncthreads();
    }
    if (threadIdx.x == 0) *alpha_out = red[0];
}

// NCCL route (no peer mailboxes, or more than MB_MAXN values): plain second stage
__global__ void __launch_bounds__(256)
finish_plain_kernel(const double *__restrict__ partial, int nblocks, int n, double *__restrict__ out)
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
        if (threadIdx.x < w) red[threadIdx.x] += red[threadIdx.x + w];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[i] = red[0];
}

__global__ void __launch_bounds__(256)
coef_plain_kernel(const double *__restrict__ T, int k, int n, const double *__restrict__ g,
                  double *__restrict__ coef)
{
    const int warp = (blockIdx.x * 256 + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp < 2 * k)
    {
        const int j = warp < k ? warp : warp - k;
        const double *gv = warp < k ? g + k : g;
        double s = 0.0;
        for (int l = lane; l < k; l += 32)
            s = fma(T[j + (size_t)l * k], gv[l], s);
        for (int o = 16; o > 0; o >>= 1)
            s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) coef[warp] = s;
    }
    else
    {
        const int i = 2 * k + (warp - 2 * k) * 32 + lane;
        if (i < n) coef[i] = g[i];
    }
}

__global__ void scale_plain_kernel(double *v, int n, double *norm_out, const double *scale_sq)
{
    // single CTA
    __shared__ double sc;
    if (threadIdx.x == 0)
    {
        if (norm_out) { *norm_out = v[n]; sc = 1.0 / sqrt(fabs(v[n])); }
        else sc = 1.0 / sqrt(fabs(*scale_sq));
    }
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x)
        v[i] *= sc;
}

}

namespace rails
{

int finish_reduction(const double *partial, int nblocks, int n, double *out, bool over_ranks,
                     const double *T, int k, double *coef, double *norm_out, const double *scale_sq,
                     double *alpha_out)
{
    Context &c = ctx();
    if (n <= 0) return RAILS_OK;
    const bool multi = over_ranks && c.comm && c.nranks > 1;
    if ((!multi || c.p2p) && n <= MB_MAXN)
    {
        Mailbox mb = c.mb;
        if (!multi) { mb.nranks = 1; mb.rank = 0; }
        unsigned long long seq = 0;
        if (multi) seq = ++c.mb_seq;
        static int configured_device = -1;
        if (configured_device != c.device)
        {
            RAILS_CUDA_TRY(cudaFuncSetAttribute(finish_exchange_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                2 * MB_MAXN * (int)sizeof(double)));
            configured_device = c.device;
        }
        RAILS_LAUNCH(finish_exchange_kernel, 1, FX_THREADS, (size_t)(alpha_out ? 2 * n : n) * sizeof(double), partial, nblocks, n, out,
                     mb, seq, c.d_mb_error, c.mb_timeout_cycles, T, k, coef, norm_out, scale_sq, alpha_out);
        RAILS_CHECK_LAUNCH();
        return RAILS_OK;
    }
    if (!(nblocks == 1 && partial == out))
    {
        RAILS_LAUNCH(finish_plain_kernel, n, 256, 0, partial, nblocks, n, out);
        RAILS_CHECK_LAUNCH();
    }
    if (multi)
        RAILS_NCCL_TRY(ncclAllReduce(out, out, (size_t)n, ncclDouble, ncclSum, c.comm, c.stream));
    if (norm_out || scale_sq)
    {
        if (norm_out) n -= 1;
        RAILS_LAUNCH(scale_plain_kernel, 1, 256, 0, out, n, norm_out, scale_sq);
        RAILS_CHECK_LAUNCH();
    }
    if (coef)
    {
        const int warps = 2 * k + (n - 2 * k + 31) / 32;
        RAILS_LAUNCH(coef_plain_kernel, (warps * 32 + 255) / 256, 256, 0, T, k, n, out, coef);
        RAILS_CHECK_LAUNCH();
        if (alpha_out)
        {
            RAILS_LAUNCH(alpha_plain_kernel, 1, 256, 0, out, coef, n, alpha_out);
            RAILS_CHECK_LAUNCH();
        }
    }
    return RAILS_OK;
}

int lanczos_coef(const double *T, int k, int n, const double *g, double *coef, double *alpha_out)
{
    const int warps = 2 * k + (n - 2 * k + 31) / 32;
    RAILS_LAUNCH(coef_plain_kernel, (warps * 32 + 255) / 256, 256, 0, T, k, n, g, coef);
    RAILS_CHECK_LAUNCH();
    if (alpha_out)
    {
        RAILS_LAUNCH(alpha_plain_kernel, 1, 256, 0, g, coef, n, alpha_out);
        RAILS_CHECK_LAUNCH();
    }
    return RAILS_OK;
}

// Called from rails_cuda_comm_init after the NCCL communicator exists: allocates
// this rank's mailbox, exchanges CUDA IPC handles through NCCL and maps the
// peers'.  Any failure leaves p2p = false (NCCL all-reduce is used instead).
int setup_mailboxes()
{
    Context &c = ctx();
    c.p2p = false;
    if (c.nranks < 2 || c.nranks > MB_MAX_RANKS)
        return RAILS_OK;
    const char *off = getenv("RAILS_NO_P2P");
    if (off && strcmp(off, "0") != 0)
        return RAILS_OK;
    const size_t flag_bytes = 4096;
    const size_t data_bytes = (size_t)MB_SLOTS * MB_MAX_RANKS * MB_MAXN * sizeof(double);
    if (cudaMalloc(&c.mb_local, flag_bytes + data_bytes) != cudaSuccess) { cudaGetLastError(); return RAILS_OK; }
    cudaMemset(c.mb_local, 0, flag_bytes + data_bytes);
    cudaIpcMemHandle_t mine;
    int ok = cudaIpcGetMemHandle(&mine, c.mb_local) == cudaSuccess ? 1 : 0;
    if (!ok) cudaGetLastError();
    // gather handles (+ ok flag) over NCCL
    const size_t rec = sizeof(cudaIpcMemHandle_t) + 8;
    char *d_all = nullptr;
    RAILS_CUDA_TRY(cudaMalloc(&d_all, rec * c.nranks));
    std::vector<char> h_all(rec * c.nranks, 0);
    memcpy(h_all.data() + rec * c.rank, &mine, sizeof(mine));
    h_all[rec * c.rank + sizeof(mine)] = (char)ok;
    RAILS_CUDA_TRY(cudaMemcpy(d_all + rec * c.rank, h_all.data() + rec * c.rank, rec, cudaMemcpyHostToDevice));
    RAILS_NCCL_TRY(ncclAllGather(d_all + rec * c.rank, d_all, rec, ncclChar, c.comm, c.stream));
    RAILS_CUDA_TRY(cudaStreamSynchronize(c.stream));
    RAILS_CUDA_TRY(cudaMemcpy(h_all.data(), d_all, rec * c.nranks, cudaMemcpyDeviceToHost));
    cudaFree(d_all);
    int all_ok = 1;
    for (int p = 0; p < c.nranks; ++p)
        if (!h_all[rec * p + sizeof(mine)]) all_ok = 0;
    if (all_ok)
        for (int p = 0; p < c.nranks && all_ok; ++p)
        {
            if (p == c.rank) { c.mb_mapped[p] = c.mb_local; continue; }
            cudaIpcMemHandle_t h;
            memcpy(&h, h_all.data() + rec * p, sizeof(h));
            if (cudaIpcOpenMemHandle(&c.mb_mapped[p], h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess)
            {
                cudaGetLastError();
                c.mb_mapped[p] = nullptr;
                all_ok = 0;
            }
        }
    // everybody must agree (a rank that failed to map a peer makes all fall back)
    int *d_ok = nullptr;
    RAILS_CUDA_TRY(cudaMalloc(&d_ok, sizeof(int)));
    RAILS_CUDA_TRY(cudaMemcpy(d_ok, &all_ok, sizeof(int), cudaMemcpyHostToDevice));
    RAILS_NCCL_TRY(ncclAllReduce(d_ok, d_ok, 1, ncclInt, ncclMin, c.comm, c.stream));
    RAILS_CUDA_TRY(cudaStreamSynchronize(c.stream));
    RAILS_CUDA_TRY(cudaMemcpy(&all_ok, d_ok, sizeof(int), cudaMemcpyDeviceToHost));
    cudaFree(d_ok);
    if (!all_ok)
    {
        fprintf(stderr, "rails_cuda: peer mailboxes unavailable, small reductions use NCCL\n");
        return RAILS_OK;
    }
    for (int p = 0; p < c.nranks; ++p)
    {
        c.mb.flags[p] = reinterpret_cast<unsigned long long *>(c.mb_mapped[p]);
        c.mb.data[p] = reinterpret_cast<double *>((char *)c.mb_mapped[p] + flag_bytes);
    }
    c.mb.rank = c.rank;
    c.mb.nranks = c.nranks;
    c.mb_seq = 0;
    c.p2p = true;
    return RAILS_OK;
}

void release_mailboxes()
{
    Context &c = ctx();
    for (int p = 0; p < MB_MAX_RANKS; ++p)
    {
        if (c.mb_mapped[p] && c.mb_mapped[p] != c.mb_local) cudaIpcCloseMemHandle(c.mb_mapped[p]);
        c.mb_mapped[p] = nullptr;
    }
    if (c.mb_local) cudaFree(c.mb_local);
    c.mb_local = nullptr;
    c.p2p = false;
}

}
