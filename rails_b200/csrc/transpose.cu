// Transposed CSR built on the device (Operator::transpose() + apply,
// /root/reference/src/Epetra_OperatorWrapper.cpp:50-55,81 -- Epetra flips a flag and
// walks the same rows; here A' is a CSR matrix of its own so that the transposed apply
// is the same staged SpMM kernel as the plain one).
//
//   histogram of the column indices -> exclusive scan = row pointers of A'
//   stable LSD radix sort of (column, position) pairs by column, 8 bits a pass
//   gather: A'.colidx[q] = row of position p_q (binary search in rowptr), A'.vals[q] = vals[p_q]
//
// The sort is STABLE and CSR positions ascend with the row, so every row of A' lists
// its entries by ascending original row whatever the thread schedule: the transposed
// apply sums in a fixed order and is bit-reproducible.  Nothing goes over PCIe but the
// per-chunk sums of the scans (a few thousand ints).
#include "internal.h"

#include <algorithm>
#include <vector>

using rails::ctx;

namespace
{

constexpr int SCAN_CHUNK = 2048;      // ints per warp in the scan kernels
constexpr int SORT_THREADS = 256;
constexpr int SORT_ROUNDS = 8;        // tile = 8 rounds x 256 threads, index order = (round, thread)
constexpr int SORT_TILE = SORT_THREADS * SORT_ROUNDS;

// ---- exclusive scan of n ints in place: chunk sums -> host scan of the sums -> in-chunk scan
__global__ void __launch_bounds__(256)
chunk_sum_kernel(const int *__restrict__ data, long long n, long long nchunks, int *__restrict__ sums)
{
    const int lane = threadIdx.x & 31;
    for (long long ch = (long long)blockIdx.x * 8 + (threadIdx.x >> 5); ch < nchunks; ch += (long long)gridDim.x * 8)
    {
        const long long b = ch * SCAN_CHUNK, e = min(n, b + SCAN_CHUNK);
        int s = 0;
        for (long long i = b + lane; i < e; i += 32) s += data[i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) sums[ch] = s;
    }
}

__global__ void __launch_bounds__(256)
chunk_scan_kernel(int *__restrict__ data, long long n, long long nchunks, const int *__restrict__ offsets)
{
    const int lane = threadIdx.x & 31;
    for (long long ch = (long long)blockIdx.x * 8 + (threadIdx.x >> 5); ch < nchunks; ch += (long long)gridDim.x * 8)
    {
        const long long b = ch * SCAN_CHUNK, e = min(n, b + SCAN_CHUNK);
        int base = offsets[ch];
        for (long long i0 = b; i0 < e; i0 += 32)
        {
            const long long i = i0 + lane;
            const int v = i < e ? data[i] : 0;
            int incl = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1)
            {
                const int t = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += t;
            }
            if (i < e) data[i] = base + incl - v;
            base += __shfl_sync(0xffffffffu, incl, 31);
        }
    }
}

// total = sum of all entries (returned for the caller's checks); data becomes its exclusive scan
int exclusive_scan(int *data, long long n, long long *total)
{
    rails::Context &c = ctx();
    if (total) *total = 0;
    if (n <= 0) return RAILS_OK;
    const long long nchunks = (n + SCAN_CHUNK - 1) / SCAN_CHUNK;
    int *sums = nullptr;
    int rc = rails::dev_alloc_t(&sums, (size_t)nchunks * sizeof(int));
    if (rc) return rc;
    const int nb = (int)std::min<long long>((nchunks + 7) / 8, (long long)c.sm_count * 8);
    RAILS_LAUNCH(chunk_sum_kernel, nb, 256, 0, (const int *)data, n, nchunks, sums);
    std::vector<int> h((size_t)nchunks);
    cudaError_t e = cudaMemcpyAsync(h.data(), sums, (size_t)nchunks * sizeof(int), cudaMemcpyDeviceToHost, c.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c.stream);
    if (e != cudaSuccess)
    {
        rails::dev_free(sums);
        return rails::fail(RAILS_ERR_CUDA, "exclusive_scan: %s", cudaGetErrorString(e));
    }
    long long run = 0;
    for (long long i = 0; i < nchunks; ++i)
    {
        const int s = h[i];
        h[i] = (int)run;
        run += s;
    }
    if (run >= (1LL << 31))
    {
        rails::dev_free(sums);
        return rails::fail(RAILS_ERR_ARG, "exclusive_scan: total %lld exceeds the int32 range", run);
    }
    e = cudaMemcpyAsync(sums, h.data(), (size_t)nchunks * sizeof(int), cudaMemcpyHostToDevice, c.stream);
    if (e == cudaSuccess)
    {
        RAILS_LAUNCH(chunk_scan_kernel, nb, 256, 0, data, n, nchunks, (const int *)sums);
        e = cudaStreamSynchronize(c.stream);        // h must outlive the copy
    }
    rails::dev_free(sums);
    if (e != cudaSuccess)
        return rails::fail(RAILS_ERR_CUDA, "exclusive_scan: %s", cudaGetErrorString(e));
    if (total) *total = run;
    return RAILS_OK;
}

// ---- one stable radix pass on bits [shift, shift + 8) --------------------------------
// hist[digit * ntiles + tile] = how many keys of the tile carry the digit
__global__ void __launch_bounds__(SORT_THREADS)
radix_hist_kernel(const int *__restrict__ keys, long long n, int shift, long long ntiles, int *__restrict__ hist)
{
    __shared__ int bins[256];
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x)
    {
        bins[threadIdx.x] = 0;
        __syncthreads();
        const long long b = tile * SORT_TILE;
#pragma unroll
        for (int r = 0; r < SORT_ROUNDS; ++r)
        {
            const long long i = b + r * SORT_THREADS + threadIdx.x;
            if (i < n) atomicAdd(&bins[((unsigned)keys[i] >> shift) & 255u], 1);
        }
        __syncthreads();
        hist[(long long)threadIdx.x * ntiles + tile] = bins[threadIdx.x];
        __syncthreads();
    }
}

// `offs` = exclusive scan of hist: where the tile's keys with each digit start.  Inside
// a tile the keys go out in index order (round by round, warp by warp, lane by lane).
__global__ void __launch_bounds__(SORT_THREADS)
radix_scatter_kernel(const int *__restrict__ keys, const int *__restrict__ vals, long long n, int shift,
                     long long ntiles, const int *__restrict__ offs, int *__restrict__ keys_out,
                     int *__restrict__ vals_out)
{
    __shared__ int base[256];
    __shared__ int wcount[SORT_THREADS / 32][256];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x)
    {
        base[threadIdx.x] = offs[(long long)threadIdx.x * ntiles + tile];
        const long long b = tile * SORT_TILE;
        for (int r = 0; r < SORT_ROUNDS; ++r)
        {
            for (int w = 0; w < SORT_THREADS / 32; ++w) wcount[w][threadIdx.x] = 0;
            __syncthreads();
            const long long i = b + r * SORT_THREADS + threadIdx.x;
            const bool live = i < n;
            const int key = live ? keys[i] : 0;
            const int d = live ? (int)(((unsigned)key >> shift) & 255u) : 256 + lane;     // dead lanes match nobody
            const unsigned peers = __match_any_sync(0xffffffffu, d);
            const int rank = __popc(peers & ((1u << lane) - 1u));
            if (live && rank == 0) wcount[warp][d] = __popc(peers);
            __syncthreads();
            if (live)
            {
                int pos = base[d] + rank;
                for (int w = 0; w < warp; ++w) pos += wcount[w][d];
                keys_out[pos] = key;
                vals_out[pos] = vals[i];
            }
            __syncthreads();
            int add = 0;
            for (int w = 0; w < SORT_THREADS / 32; ++w) add += wcount[w][threadIdx.x];
            base[threadIdx.x] += add;
            __syncthreads();
        }
    }
}

__global__ void __launch_bounds__(256)
iota_kernel(int *__restrict__ p, long long n)
{
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256)
        p[i] = (int)i;
}

__global__ void __launch_bounds__(256)
count_columns_kernel(const int *__restrict__ colidx, long long nnz, int *__restrict__ counts)
{
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < nnz; i += (long long)gridDim.x * 256)
        atomicAdd(&counts[colidx[i]], 1);          // integer adds: the result does not depend on the order
}

// position p of the original matrix -> its row: the last r with rowptr[r] <= p
__global__ void __launch_bounds__(256)
gather_transposed_kernel(const int *__restrict__ pos, long long nnz, const int *__restrict__ rowptr, long long rows,
                         const double *__restrict__ vals, int *__restrict__ colidx_t, double *__restrict__ vals_t)
{
    for (long long q = (long long)blockIdx.x * 256 + threadIdx.x; q < nnz; q += (long long)gridDim.x * 256)
    {
        const int p = pos[q];
        long long a = 0, b = rows;              // invariant: rowptr[a] <= p < rowptr[b]
        while (b - a > 1)
        {
            const long long mid = (a + b) >> 1;
            if (rowptr[mid] <= p) a = mid; else b = mid;
        }
        colidx_t[q] = (int)a;
        vals_t[q] = vals[p];
    }
}

}

namespace rails
{

// Device arrays of the CSR form of A' for an A of `rows` x `ncols` with `nnz` entries
// (all device pointers).  The three outputs come from dev_alloc (caller frees); sized
// ncols + 1, nnz (+ 64 bytes of slack for the staged kernel's rounded-up bulk copies).
int csr_transpose_device(long long rows, long long ncols, long long nnz, const int *rowptr, const int *colidx,
                         const double *vals, int **rowptr_t, int **colidx_t, double **vals_t)
{
    Context &c = ctx();
    *rowptr_t = nullptr; *colidx_t = nullptr; *vals_t = nullptr;
    int rc = dev_alloc_t(rowptr_t, (size_t)(ncols + 1) * sizeof(int));
    if (!rc) rc = dev_alloc_t(colidx_t, (size_t)std::max<long long>(nnz, 1) * sizeof(int) + 64);
    if (!rc) rc = dev_alloc_t(vals_t, (size_t)std::max<long long>(nnz, 1) * sizeof(double) + 64);
    int *k0 = nullptr, *k1 = nullptr, *p0 = nullptr, *p1 = nullptr, *hist = nullptr;
    auto cleanup = [&](bool all) {
        dev_free(k0); dev_free(k1); dev_free(p0); dev_free(p1); dev_free(hist);
        if (all) { dev_free(*rowptr_t); dev_free(*colidx_t); dev_free(*vals_t); *rowptr_t = *colidx_t = nullptr; *vals_t = nullptr; }
    };
    if (rc) { cleanup(true); return rc; }
    RAILS_CUDA_TRY(cudaMemsetAsync(*rowptr_t, 0, (size_t)(ncols + 1) * sizeof(int), c.stream));
    if (nnz == 0)
        return RAILS_OK;
    const int nb = (int)std::min<long long>((nnz + 1023) / 1024, (long long)c.sm_count * 8);
    // row pointers of A': counts of each column, scanned
    RAILS_LAUNCH(count_columns_kernel, nb, 256, 0, colidx, nnz, *rowptr_t);
    RAILS_CHECK_LAUNCH();
    long long total = 0;
    rc = exclusive_scan(*rowptr_t, ncols + 1, &total);
    if (!rc && total != nnz)
        rc = fail(RAILS_ERR_STATE, "csr_transpose: %lld entries counted, %lld expected", total, nnz);
    // stable sort of (column, position) by column
    const size_t ib = (size_t)nnz * sizeof(int);
    const long long ntiles = (nnz + SORT_TILE - 1) / SORT_TILE;
    if (!rc) rc = dev_alloc_t(&k0, ib);
    if (!rc) rc = dev_alloc_t(&k1, ib);
    if (!rc) rc = dev_alloc_t(&p0, ib);
    if (!rc) rc = dev_alloc_t(&p1, ib);
    if (!rc) rc = dev_alloc_t(&hist, (size_t)256 * ntiles * sizeof(int));
    if (rc) { cleanup(true); return rc; }
    RAILS_CUDA_TRY(cudaMemcpyAsync(k0, colidx, ib, cudaMemcpyDeviceToDevice, c.stream));
    RAILS_LAUNCH(iota_kernel, nb, 256, 0, p0, nnz);
    int bits = 1;
    while (bits < 31 && (1LL << bits) < ncols) ++bits;
    const int ng = (int)std::min<long long>(ntiles, (long long)c.sm_count * 8);
    for (int shift = 0; shift < bits; shift += 8)
    {
        RAILS_LAUNCH(radix_hist_kernel, ng, SORT_THREADS, 0, (const int *)k0, nnz, shift, ntiles, hist);
        RAILS_CHECK_LAUNCH();
        rc = exclusive_scan(hist, 256 * ntiles, nullptr);
        if (rc) { cleanup(true); return rc; }
        RAILS_LAUNCH(radix_scatter_kernel, ng, SORT_THREADS, 0, (const int *)k0, (const int *)p0, nnz, shift, ntiles,
                     (const int *)hist, k1, p1);
        RAILS_CHECK_LAUNCH();
        std::swap(k0, k1);
        std::swap(p0, p1);
    }
    RAILS_LAUNCH(gather_transposed_kernel, nb, 256, 0, (const int *)p0, nnz, rowptr, rows, vals, *colidx_t, *vals_t);
    RAILS_CHECK_LAUNCH();
    RAILS_CUDA_TRY(cudaStreamSynchronize(c.stream));
    cleanup(false);
    return RAILS_OK;
}

}
