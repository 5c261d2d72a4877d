// CSR x tall-skinny multivector:  Y = op(A) X.
//
// Replaces EpetraOp::operator* (src/Epetra_OperatorWrapper.cpp:75-91 ->
// Epetra_CrsMatrix::Apply, i.e. ghost import + row loop) and the dense
// stand-in StlWrapper::operator* (src/StlWrapper.cpp:168-187) at
// src/LyapunovSolver.hpp:146.  Algorithmic traffic:
// 12 nnz + 4 (m+1) + 16 m e bytes; 1/6 flop per byte per column -> HBM bound.
//
// Kernel: rows are cut on the host into blocks of <= 256 rows whose non-zeros
// fit a shared-memory stage (sized for the matrix at creation, <= 6144 nnz).
// Persistent CTAs (SMs x resident CTAs) walk the blocks; one thread moves the
// NEXT block's values and column indices into the other stage with two 1-D
// bulk copies (cp.async.bulk + mbarrier, the TMA engine) while the CTA computes
// the current block, so the CSR arrays are read from HBM exactly once per call,
// whatever the number of columns e, without passing through registers.  Each
// thread then walks ONE row sequentially out of shared memory for EC columns
// of X at a time.  Row sums are therefore accumulated in CSR order with a
// separately rounded multiply and add, which makes the result bit-identical
// to the scalar CPU loop y_i = sum_j a_ij x_j of the oracle.  Gathers from X
// go through the read-only path; for stencil matrices neighbouring rows hit
// the same lines and the rest is served by L2.
//
// Row partition (multi-GPU): columns are renumbered [owned | ghost] at
// creation; ghost rows of X are packed by the owners and moved with grouped
// ncclSend/ncclRecv over NVLink before the kernel reads them.
#include "internal.h"

#include <algorithm>
#include <cstring>

using rails::ctx;

namespace
{

constexpr int SPMM_THREADS = 256;       // rows per block (one thread per row)
constexpr int SPMM_STAGES = 2;
constexpr int SPMM_NNZ_CAP_MAX = 6144;  // largest staging buffer (nnz per block)

// ---- mbarrier / bulk-copy (TMA 1-D) primitives --------------------------
__device__ __forceinline__ unsigned smem_u32(const void *p)
{
    return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(unsigned long long *bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity)
        : "memory");
}
// global -> shared bulk copy, completion counted in bytes on `bar`
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, unsigned long long *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// Persistent CTAs walk the row blocks cyclically (block b, b + grid, ...), so
// that at any moment the whole GPU works on one narrow band of rows and the
// gathers from X are served by L2.  One thread issues the two bulk copies of
// the NEXT block (values, column indices) into the other stage while the CTA
// computes the current one: the CSR arrays never pass through registers and
// every SM keeps tens of KB of loads in flight.
//
// Shared memory per stage: cap + 2 doubles, then cap + 4 ints.  Bulk copies
// need 16-byte aligned addresses and sizes, so a copy starts at the aligned
// element at or before p0 (the arrays carry 64 bytes of slack at the end).
template<int EC, bool GHOST>
__global__ void __launch_bounds__(SPMM_THREADS)
spmm_kernel(const int *__restrict__ rowptr, const int *__restrict__ colidx,
            const double *__restrict__ vals, const int *__restrict__ blk_row, int nblk, int cap,
            const double *__restrict__ X, long long ldx, long long nlocal,
            const double *__restrict__ G, long long ldg,
            double *__restrict__ Y, long long ldy, int ncols)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ unsigned long long full[SPMM_STAGES];
    __shared__ double red[SPMM_THREADS / 32];

    const size_t v_bytes = (size_t)(cap + 2) * sizeof(double);
    const size_t stage_bytes = v_bytes + (size_t)(cap + 4) * sizeof(int);   // multiple of 16
    const int tid = threadIdx.x;

    if (tid == 0)
    {
        for (int s = 0; s < SPMM_STAGES; ++s)
            mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    auto issue = [&](int b, int s) {
        const int r0 = blk_row[b], r1 = blk_row[b + 1];
        const int p0 = rowptr[r0], p1 = rowptr[r1];
        if (p1 - p0 > cap || p1 == p0)
        {
            mbar_arrive(&full[s]);      // nothing staged: a long single row, or an empty block
            return;
        }
        unsigned char *base = smem_raw + (size_t)s * stage_bytes;
        const int pv = p0 & ~1, pc = p0 & ~3;
        const unsigned bv = (unsigned)(((p1 - pv) * 8 + 15) & ~15);
        const unsigned bc = (unsigned)(((p1 - pc) * 4 + 15) & ~15);
        mbar_expect_tx(&full[s], bv + bc);
        bulk_g2s(base, vals + pv, bv, &full[s]);
        bulk_g2s(base + v_bytes, colidx + pc, bc, &full[s]);
    };

    int blk = blockIdx.x;
    if (tid == 0 && blk < nblk)
        issue(blk, 0);

    for (int it = 0; blk < nblk; blk += gridDim.x, ++it)
    {
        const int s = it & (SPMM_STAGES - 1);
        if (tid == 0 && blk + (int)gridDim.x < nblk)
            issue(blk + gridDim.x, s ^ 1);

        const int r0 = blk_row[blk], r1 = blk_row[blk + 1];
        const int p0 = rowptr[r0], p1 = rowptr[r1];
        const int row = r0 + tid;
        int a = 0, e = 0;
        if (row < r1)
        {
            a = rowptr[row];
            e = rowptr[row + 1];
        }
        mbar_wait(&full[s], (unsigned)(it / SPMM_STAGES) & 1u);

        if (p1 - p0 <= cap)
        {
            const unsigned char *base = smem_raw + (size_t)s * stage_bytes;
            const double *sv = reinterpret_cast<const double *>(base) - (p0 & ~1);
            const int *sc = reinterpret_cast<const int *>(base + v_bytes) - (p0 & ~3);
            if (row < r1)
            {
                for (int c0 = 0; c0 < ncols; c0 += EC)
                {
                    double acc[EC];
#pragma unroll
                    for (int t = 0; t < EC; ++t)
                        acc[t] = 0.0;
#pragma unroll 4
                    for (int p = a; p < e; ++p)
                    {
                        const int col = sc[p];
                        const double v = sv[p];
                        const double *xp = X + col;
                        long long ld = ldx;
                        if (GHOST && col >= nlocal) { xp = G + (col - nlocal); ld = ldg; }
                        xp += (size_t)c0 * ld;
#pragma unroll
                        for (int t = 0; t < EC; ++t)
                            if (EC == 1 || c0 + t < ncols)
                                acc[t] = __dadd_rn(acc[t], __dmul_rn(v, __ldg(xp + (size_t)t * ld)));
                    }
#pragma unroll
                    for (int t = 0; t < EC; ++t)
                        if (EC == 1 || c0 + t < ncols)
                            Y[row + (size_t)(c0 + t) * ldy] = acc[t];
                }
            }
        }
        else
        {
            // a single row longer than the staging buffer: whole CTA on one row
            for (int c = 0; c < ncols; ++c)
            {
                double sum = 0.0;
                for (int p = p0 + tid; p < p1; p += SPMM_THREADS)
                {
                    const int col = __ldg(colidx + p);
                    const double x = (!GHOST || col < nlocal) ? __ldg(X + col + (size_t)c * ldx)
                                                              : __ldg(G + (col - nlocal) + (size_t)c * ldg);
                    sum = fma(__ldg(vals + p), x, sum);
                }
                for (int off = 16; off > 0; off >>= 1)
                    sum += __shfl_xor_sync(0xffffffffu, sum, off);
                if ((tid & 31) == 0) red[tid >> 5] = sum;
                __syncthreads();
                if (tid == 0)
                {
                    double tot = 0.0;
                    for (int w = 0; w < SPMM_THREADS / 32; ++w) tot += red[w];
                    Y[r0 + (size_t)c * ldy] = tot;
                }
                __syncthreads();
            }
        }
        __syncthreads();        // every thread is done with stage s before it is refilled
    }
}

// flag = 1 if any column index lies outside [0, ncols)
__global__ void __launch_bounds__(256)
check_columns_kernel(const int *__restrict__ colidx, long long nnz, int ncols, int *flag)
{
    bool bad = false;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < nnz; i += (long long)gridDim.x * 256)
    {
        const int c = colidx[i];
        bad |= (c < 0) | (c >= ncols);
    }
    if (__syncthreads_or(bad) && threadIdx.x == 0)
        atomicOr(flag, 1);
}

__global__ void __launch_bounds__(256)
pack_rows_kernel(const int *__restrict__ rows, int n, const double *__restrict__ X, long long ldx,
                 double *__restrict__ buf, long long ldb)
{
    const int c = blockIdx.y;
    for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256)
        buf[i + (size_t)c * ldb] = X[rows[i] + (size_t)c * ldx];
}

void free_csr(rails_csr_s *A)
{
    if (!A) return;
    if (A->transposed) free_csr(A->transposed);
    if (ctx().ready)
    {
        rails::dev_free(A->rowptr); rails::dev_free(A->colidx); rails::dev_free(A->vals);
        rails::dev_free(A->blk_row);
        rails::dev_free(A->send_rows);
        rails::dev_free(A->send_buf);
        rails::dev_free(A->ghost_buf);
        ctx().bytes_in_use.fetch_sub((long long)A->bytes);
    }
    delete A;
}

// uploads CSR arrays whose column indices are already local ([owned|ghost])
// cuts the rows into blocks for the staged kernel (host row pointers in) and uploads
// the block table
int plan_blocks(rails_csr_s *A, const int *rowptr)
{
    const long long rows = A->rows;
    // staging capacity: what a full block of SPMM_THREADS rows needs, bounded
    int cap = 256;
    for (long long r = 0; r < rows; r += SPMM_THREADS)
        cap = std::max(cap, rowptr[std::min(rows, r + SPMM_THREADS)] - rowptr[r]);
    cap = std::min((cap + 63) & ~63, SPMM_NNZ_CAP_MAX);
    A->blk_cap = cap;
    std::vector<int> blk;
    blk.push_back(0);
    int maxnnz = 0;
    long long r = 0;
    while (r < rows)
    {
        long long end = std::min(rows, r + SPMM_THREADS);
        // shrink until the block fits the staging buffer (keep >= 1 row)
        while (end > r + 1 && rowptr[end] - rowptr[r] > cap)
            --end;
        maxnnz = std::max(maxnnz, rowptr[end] - rowptr[r]);
        blk.push_back((int)end);
        r = end;
    }
    A->nblk = (int)blk.size() - 1;
    A->max_blk_nnz = maxnnz;
    const size_t b_blk = blk.size() * sizeof(int);
    int rc = rails::dev_alloc_t(&A->blk_row, b_blk);
    if (rc) return rc;
    A->bytes += b_blk;
    ctx().bytes_in_use.fetch_add((long long)b_blk);
    return rails::copy_h2d(A->blk_row, blk.data(), b_blk);
}

int upload_csr(rails_csr_s *A, const int *rowptr, const int *colidx, const double *vals, bool validate = true)
{
    const long long rows = A->rows, nnz = A->nnz;
    int rca = plan_blocks(A, rowptr);
    if (rca) return rca;
    // 64 bytes of slack: bulk copies are rounded up to 16 bytes
    size_t b_rp = (size_t)(rows + 1) * sizeof(int), b_ci = (size_t)nnz * sizeof(int) + 64,
           b_v = (size_t)nnz * sizeof(double) + 64;
    rca = rails::dev_alloc_t(&A->rowptr, b_rp);
    if (!rca) rca = rails::dev_alloc_t(&A->colidx, b_ci);
    if (!rca) rca = rails::dev_alloc_t(&A->vals, b_v);
    if (rca) return rca;
    A->bytes += b_rp + b_ci + b_v;
    ctx().bytes_in_use.fetch_add((long long)(b_rp + b_ci + b_v));
    int rc = rails::copy_h2d(A->rowptr, rowptr, b_rp);
    if (rc) return rc;
    if (nnz)
    {
        rc = rails::copy_h2d(A->colidx, colidx, (size_t)nnz * sizeof(int));
        if (rc) return rc;
        rc = rails::copy_h2d(A->vals, vals, (size_t)nnz * sizeof(double));
        if (rc) return rc;
    }
    // column indices are validated where they now live: one streaming pass on
    // the device instead of a scalar host loop over nnz entries
    if (nnz && validate)
    {
        rc = rails::ensure_small(1);
        if (rc) return rc;
        int *flag = reinterpret_cast<int *>(ctx().d_small);
        RAILS_CUDA_TRY(cudaMemsetAsync(flag, 0, sizeof(int), ctx().stream));
        const int nb = (int)std::min<long long>((nnz + 1023) / 1024, (long long)ctx().sm_count * 8);
        RAILS_LAUNCH(check_columns_kernel, nb, 256, 0, A->colidx, nnz, (int)(A->ncols_local + A->nghost), flag);
        RAILS_CHECK_LAUNCH();
        int bad = 0;
        RAILS_CUDA_TRY(cudaMemcpyAsync(&bad, flag, sizeof(int), cudaMemcpyDeviceToHost, ctx().stream));
        RAILS_CUDA_TRY(cudaStreamSynchronize(ctx().stream));
        if (bad)
            return rails::fail(RAILS_ERR_ARG, "csr_create: column index out of range");
    }
    return RAILS_OK;
}

int owner_of(long long gcol, long long m, int P)
{
    long long base = m / P, rem = m % P;
    long long split = rem * (base + 1);
    if (gcol < split) return (int)(gcol / (base + 1));
    return (int)(rem + (gcol - split) / std::max<long long>(base, 1));
}

// ---- halo plan on the device ------------------------------------------------------
// The same plan as rails_csr_halo_plan (the host restatement the CPU tests pin), built
// where the column indices already are: mark the ghost columns in a byte map of the
// global index range, compact the map in ascending order (chunk counts -> host scan of
// a few thousand counters -> ordered write), renumber in place.  A 58 M-entry rank
// block takes a few hundred microseconds instead of ~0.4 s of host loops.
constexpr int GH_CHUNK = 2048;          // flags per warp

__global__ void __launch_bounds__(256)
mark_ghost_kernel(const int *__restrict__ colidx, long long nnz, long long global_rows, long long lo, long long hi,
                  unsigned char *__restrict__ flags, int *bad)
{
    bool b = false;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < nnz; i += (long long)gridDim.x * 256)
    {
        const long long g = colidx[i];
        if (g < 0 || g >= global_rows)
            b = true;
        else if (g < lo || g >= hi)
            flags[g] = 1;
    }
    if (__syncthreads_or(b) && threadIdx.x == 0)
        atomicOr(bad, 1);
}

__global__ void __launch_bounds__(256)
count_ghost_kernel(const unsigned char *__restrict__ flags, long long nchunks, int *__restrict__ counts)
{
    const int lane = threadIdx.x & 31;
    for (long long ch = (long long)blockIdx.x * 8 + (threadIdx.x >> 5); ch < nchunks; ch += (long long)gridDim.x * 8)
    {
        const unsigned *w = reinterpret_cast<const unsigned *>(flags + ch * GH_CHUNK);
        int c = 0;
#pragma unroll
        for (int it = 0; it < GH_CHUNK / 128; ++it)
            c += __popc(w[it * 32 + lane] & 0x01010101u);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
            c += __shfl_xor_sync(0xffffffffu, c, o);
        if (lane == 0) counts[ch] = c;
    }
}

// list[offsets[ch] ..] = the flagged indices of chunk ch in ascending order
__global__ void __launch_bounds__(256)
write_ghost_kernel(const unsigned char *__restrict__ flags, long long nchunks, const int *__restrict__ offsets,
                   int *__restrict__ list)
{
    const int lane = threadIdx.x & 31;
    for (long long ch = (long long)blockIdx.x * 8 + (threadIdx.x >> 5); ch < nchunks; ch += (long long)gridDim.x * 8)
    {
        const unsigned *w = reinterpret_cast<const unsigned *>(flags + ch * GH_CHUNK);
        int base = offsets[ch];
        if (offsets[ch + 1] == base) continue;
        for (int it = 0; it < GH_CHUNK / 128; ++it)
        {
            const unsigned f = w[it * 32 + lane] & 0x01010101u;
            const int c = __popc(f);
            int incl = c;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1)
            {
                const int t = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += t;
            }
            int pos = base + incl - c;
            const long long i0 = ch * GH_CHUNK + (long long)it * 128 + lane * 4;
#pragma unroll
            for (int b = 0; b < 4; ++b)
                if (f & (1u << (8 * b)))
                    list[pos++] = (int)(i0 + b);
            base += __shfl_sync(0xffffffffu, incl, 31);
        }
    }
}

// owned columns shift down; a ghost column becomes rows + its position in the sorted
// ghost list (binary search: ghost entries are the few that cross the partition)
__global__ void __launch_bounds__(256)
renumber_kernel(int *__restrict__ colidx, long long nnz, long long lo, long long hi, const int *__restrict__ list,
                int nlist)
{
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < nnz; i += (long long)gridDim.x * 256)
    {
        const long long g = colidx[i];
        if (g >= lo && g < hi)
            colidx[i] = (int)(g - lo);
        else
        {
            int a = 0, b = nlist;
            while (a < b)
            {
                const int mid = (a + b) >> 1;
                if (list[mid] < (int)g) a = mid + 1; else b = mid;
            }
            colidx[i] = (int)(hi - lo) + a;
        }
    }
}

// A->colidx holds GLOBAL columns on entry and [owned | ghost] local ones on return;
// ghosts = the ghost rows in ascending global order (host copy).
int device_halo_plan(int *colidx_dev, long long nnz, long long rows, long long gm, long long row_offset,
                     std::vector<int> &ghosts)
{
    rails::Context &c = ctx();
    const long long lo = row_offset, hi = row_offset + rows;
    ghosts.clear();
    if (nnz == 0) return RAILS_OK;
    const long long nchunks = (gm + GH_CHUNK - 1) / GH_CHUNK;
    // scratch from the block cache: the next operator of the same shape reuses it
    unsigned char *flags = nullptr;
    int *counts = nullptr, *list = nullptr;
    int rc = rails::ensure_small(1);
    if (rc) return rc;
    int *bad = reinterpret_cast<int *>(c.d_small);
    auto cleanup = [&]() { rails::dev_free(flags); rails::dev_free(counts); rails::dev_free(list); };
    rc = rails::dev_alloc_t(&flags, (size_t)nchunks * GH_CHUNK);
    if (!rc) rc = rails::dev_alloc_t(&counts, (size_t)(nchunks + 1) * sizeof(int));
    if (rc) { cleanup(); return rc; }
    RAILS_CUDA_TRY(cudaMemsetAsync(flags, 0, (size_t)nchunks * GH_CHUNK, c.stream));
    RAILS_CUDA_TRY(cudaMemsetAsync(bad, 0, sizeof(int), c.stream));
    const int nb = (int)std::min<long long>((nnz + 1023) / 1024, (long long)c.sm_count * 8);
    RAILS_LAUNCH(mark_ghost_kernel, nb, 256, 0, (const int *)colidx_dev, nnz, gm, lo, hi, flags, bad);
    const int nbc = (int)std::min<long long>((nchunks + 7) / 8, (long long)c.sm_count * 8);
    RAILS_LAUNCH(count_ghost_kernel, nbc, 256, 0, (const unsigned char *)flags, nchunks, counts);
    RAILS_CHECK_LAUNCH();
    std::vector<int> h(nchunks + 1);
    int hbad = 0;
    RAILS_CUDA_TRY(cudaMemcpyAsync(h.data(), counts, (size_t)nchunks * sizeof(int), cudaMemcpyDeviceToHost, c.stream));
    RAILS_CUDA_TRY(cudaMemcpyAsync(&hbad, bad, sizeof(int), cudaMemcpyDeviceToHost, c.stream));
    RAILS_CUDA_TRY(cudaStreamSynchronize(c.stream));
    if (hbad)
    {
        cleanup();
        return rails::fail(RAILS_ERR_ARG, "halo_plan: column out of range");
    }
    long long total = 0;
    for (long long i = 0; i < nchunks; ++i)
    {
        const int n = h[i];
        h[i] = (int)total;
        total += n;
    }
    h[nchunks] = (int)total;
    if (total + rows >= (1LL << 31))
    {
        cleanup();
        return rails::fail(RAILS_ERR_ARG, "csr_create: owned + ghost columns exceed the int32 index range");
    }
    rc = rails::dev_alloc_t(&list, std::max<size_t>((size_t)total, 1) * sizeof(int));
    if (rc) { cleanup(); return rc; }
    RAILS_CUDA_TRY(cudaMemcpyAsync(counts, h.data(), (size_t)(nchunks + 1) * sizeof(int), cudaMemcpyHostToDevice, c.stream));
    RAILS_LAUNCH(write_ghost_kernel, nbc, 256, 0, (const unsigned char *)flags, nchunks, (const int *)counts, list);
    RAILS_LAUNCH(renumber_kernel, nb, 256, 0, colidx_dev, nnz, lo, hi, (const int *)list, (int)total);
    RAILS_CHECK_LAUNCH();
    ghosts.resize((size_t)total);
    if (total)
        RAILS_CUDA_TRY(cudaMemcpyAsync(ghosts.data(), list, (size_t)total * sizeof(int), cudaMemcpyDeviceToHost, c.stream));
    RAILS_CUDA_TRY(cudaStreamSynchronize(c.stream));
    cleanup();
    return RAILS_OK;
}

// owners of the ghost rows (ascending global order = grouped by owner in rank order):
// how many come from each rank, and their row ids in the owner's numbering
void ghost_owners(std::vector<int> const &ghosts, long long global_rows, int P, int *need_from, int *owner_row)
{
    for (int p = 0; p < P; ++p) need_from[p] = 0;
    for (size_t i = 0; i < ghosts.size(); ++i)
    {
        const int p = owner_of(ghosts[i], global_rows, P);
        need_from[p]++;
        long long b, e;
        rails_partition_rows(global_rows, P, p, &b, &e);
        if (owner_row) owner_row[i] = (int)(ghosts[i] - b);
    }
}

int ensure_halo_cols(rails_csr_s *A, int ncols)
{
    if (ncols <= A->halo_cols_cap) return RAILS_OK;
    rails::dev_free(A->send_buf); A->send_buf = nullptr;
    rails::dev_free(A->ghost_buf); A->ghost_buf = nullptr;
    int rc = RAILS_OK;
    if (A->nsend) rc = rails::dev_alloc_t(&A->send_buf, (size_t)A->nsend * ncols * sizeof(double));
    if (!rc && A->nghost) rc = rails::dev_alloc_t(&A->ghost_buf, (size_t)A->nghost * ncols * sizeof(double));
    if (rc) return rc;
    A->halo_cols_cap = ncols;
    return RAILS_OK;
}

int halo_exchange(rails_csr_s *A, const double *X, long long ldx, int ncols)
{
    rails::Context &c = ctx();
    int rc = ensure_halo_cols(A, ncols);
    if (rc) return rc;
    if (A->nsend)
    {
        dim3 grid(std::min((A->nsend + 255) / 256, c.sm_count * 4), ncols);
        RAILS_LAUNCH(pack_rows_kernel, grid, 256, 0, A->send_rows, A->nsend, X, ldx, A->send_buf,
                     (long long)A->nsend);
        RAILS_CHECK_LAUNCH();
    }
    RAILS_NCCL_TRY(ncclGroupStart());
    long long soff = 0, goff = 0;
    for (int p = 0; p < c.nranks; ++p)
    {
        for (int col = 0; col < ncols; ++col)
        {
            if (A->send_count[p])
                RAILS_NCCL_TRY(ncclSend(A->send_buf + soff + (size_t)col * A->nsend, A->send_count[p],
                                        ncclDouble, p, c.comm, c.stream));
            if (A->recv_count[p])
                RAILS_NCCL_TRY(ncclRecv(A->ghost_buf + goff + (size_t)col * A->nghost, A->recv_count[p],
                                        ncclDouble, p, c.comm, c.stream));
        }
        soff += A->send_count[p];
        goff += A->recv_count[p];
    }
    RAILS_NCCL_TRY(ncclGroupEnd());
    return RAILS_OK;
}

int launch_spmm(rails_csr_s *A, const double *X, long long ldx, int ncols, double *Y, long long ldy)
{
    if (A->rows == 0 || ncols == 0) return RAILS_OK;
    const double *G = A->ghost_buf;
    const long long ldg = A->nghost;
    const size_t smem = SPMM_STAGES * ((size_t)(A->blk_cap + 2) * sizeof(double) + (size_t)(A->blk_cap + 4) * sizeof(int));
#define RAILS_SPMM_CASE(EC, GH)                                                                      \
    do {                                                                                             \
        static int per_sm = 0;                                                                       \
        static size_t attr = 0, sized_for = ~(size_t)0;                                              \
        if (smem > attr)                                                                             \
        {                                                                                            \
            RAILS_CUDA_TRY(cudaFuncSetAttribute(spmm_kernel<EC, GH>,                                 \
                                                cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
            attr = smem;                                                                             \
        }                                                                                            \
        if (smem != sized_for)                                                                       \
        {                                                                                            \
            RAILS_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, spmm_kernel<EC, GH>, \
                                                                         SPMM_THREADS, smem));       \
            sized_for = smem;                                                                        \
        }                                                                                            \
        const int grid = std::min(A->nblk, std::max(per_sm, 1) * rails::ctx().sm_count);             \
        RAILS_LAUNCH((spmm_kernel<EC, GH>), grid, SPMM_THREADS, smem, A->rowptr, A->colidx, A->vals, \
                     A->blk_row, A->nblk, A->blk_cap, X, ldx, A->ncols_local, G, ldg, Y, ldy, ncols); \
    } while (0)
    const bool gh = A->nghost > 0;
    if (ncols >= 4)      { if (gh) RAILS_SPMM_CASE(4, true); else RAILS_SPMM_CASE(4, false); }
    else if (ncols >= 2) { if (gh) RAILS_SPMM_CASE(2, true); else RAILS_SPMM_CASE(2, false); }
    else                 { if (gh) RAILS_SPMM_CASE(1, true); else RAILS_SPMM_CASE(1, false); }
#undef RAILS_SPMM_CASE
    RAILS_CHECK_LAUNCH();
    return RAILS_OK;
}

}

// Host-only part of the halo plan (no CUDA call; tested on CPU with gloo):
// renumbers the global column indices of this rank's rows to [owned | ghost],
// lists the ghost rows in ascending global order (so they are grouped by owner
// in rank order), their row ids in the owner's local numbering, and how many
// are needed from each rank.  Arrays of nnz entries are enough for the outputs.
extern "C" int rails_csr_halo_plan(long long rows, long long global_rows, long long row_offset, int nranks,
                                   const int *rowptr, const int *colidx, int *local_col, int *nghost,
                                   int *ghost_global, int *ghost_owner_row, int *need_from)
{
    if (rows < 0 || nranks < 1 || !rowptr || !local_col || !nghost || !need_from)
        return rails::fail(RAILS_ERR_ARG, "halo_plan: bad arguments");
    const long long nnz = rowptr[rows];
    std::vector<int> ghosts;
    for (long long p = 0; p < nnz; ++p)
    {
        long long g = colidx[p];
        if (g < 0 || g >= global_rows)
            return rails::fail(RAILS_ERR_ARG, "halo_plan: column %lld out of range", g);
        if (g < row_offset || g >= row_offset + rows)
            ghosts.push_back((int)g);
    }
    std::sort(ghosts.begin(), ghosts.end());
    ghosts.erase(std::unique(ghosts.begin(), ghosts.end()), ghosts.end());
    *nghost = (int)ghosts.size();
    for (long long p = 0; p < nnz; ++p)
    {
        long long g = colidx[p];
        if (g >= row_offset && g < row_offset + rows)
            local_col[p] = (int)(g - row_offset);
        else
            local_col[p] = (int)(rows + (std::lower_bound(ghosts.begin(), ghosts.end(), (int)g) - ghosts.begin()));
    }
    for (int p = 0; p < nranks; ++p) need_from[p] = 0;
    for (size_t i = 0; i < ghosts.size(); ++i)
    {
        int p = owner_of(ghosts[i], global_rows, nranks);
        need_from[p]++;
        long long b, e;
        rails_partition_rows(global_rows, nranks, p, &b, &e);
        if (ghost_global) ghost_global[i] = ghosts[i];
        if (ghost_owner_row) ghost_owner_row[i] = (int)(ghosts[i] - b);
    }
    return RAILS_OK;
}

// The plan rails_csr_create builds on the device, with the signature of the host
// restatement above (tests compare the two; no communicator needed).
extern "C" int rails_csr_halo_plan_device(long long rows, long long global_rows, long long row_offset, int nranks,
                                          const int *rowptr, const int *colidx, int *local_col, int *nghost,
                                          int *ghost_global, int *ghost_owner_row, int *need_from)
{
    RAILS_REQUIRE_READY();
    if (rows < 0 || nranks < 1 || !rowptr || !local_col || !nghost || !need_from)
        return rails::fail(RAILS_ERR_ARG, "halo_plan: bad arguments");
    const long long nnz = rowptr[rows];
    int *d = nullptr;
    std::vector<int> ghosts;
    if (nnz)
    {
        RAILS_CUDA_TRY(cudaMalloc(&d, (size_t)nnz * sizeof(int)));
        int rc = rails::copy_h2d(d, colidx, (size_t)nnz * sizeof(int));
        if (!rc) rc = device_halo_plan(d, nnz, rows, global_rows, row_offset, ghosts);
        if (!rc && cudaMemcpy(local_col, d, (size_t)nnz * sizeof(int), cudaMemcpyDeviceToHost) != cudaSuccess)
            rc = rails::fail(RAILS_ERR_CUDA, "halo_plan: read-back failed");
        cudaFree(d);
        if (rc) return rc;
    }
    *nghost = (int)ghosts.size();
    ghost_owners(ghosts, global_rows, nranks, need_from, ghost_owner_row);
    if (ghost_global)
        std::copy(ghosts.begin(), ghosts.end(), ghost_global);
    return RAILS_OK;
}

extern "C" int rails_csr_create(long long rows, long long global_rows, long long global_cols,
                                long long row_offset, const int *rowptr, const int *colidx,
                                const double *vals, rails_csr_t *out)
{
    RAILS_REQUIRE_READY();
    rails::Context &c = ctx();
    if (rows < 0 || !rowptr || !out || rows >= (1LL << 31) || global_cols >= (1LL << 31))
        return rails::fail(RAILS_ERR_ARG, "csr_create: bad sizes");
    const long long nnz = rowptr[rows];
    if (nnz < 0 || (nnz > 0 && (!colidx || !vals)))
        return rails::fail(RAILS_ERR_ARG, "csr_create: bad arrays");
    const bool dist = c.nranks > 1 && rows != global_rows;
    if (dist && global_rows != global_cols)
        return rails::fail(RAILS_ERR_ARG, "csr_create: a row-partitioned operator must be square");

    rails_csr_s *A = new rails_csr_s();
    A->rows = rows; A->global_rows = global_rows; A->global_cols = global_cols;
    A->row_offset = row_offset; A->nnz = nnz;
    A->rowptr = A->colidx = A->blk_row = nullptr; A->vals = nullptr;
    A->send_rows = nullptr; A->send_buf = A->ghost_buf = nullptr;
    A->nghost = A->nsend = 0; A->halo_cols_cap = 0; A->transposed = nullptr; A->bytes = 0;
    A->partitioned = dist;
    A->recv_count.assign(c.nranks, 0);
    A->send_count.assign(c.nranks, 0);

    int rc;
    if (!dist)
    {
        A->ncols_local = global_cols;
        rc = upload_csr(A, rowptr, colidx, vals);
        if (rc) { free_csr(A); return rc; }
        *out = A;
        return RAILS_OK;
    }

    // ---- row-partitioned: renumber columns [owned | ghost], build halo plan
    A->ncols_local = rows;
    const int P = c.nranks;
    // global columns go up as they are; the plan is built and applied on the device
    rc = upload_csr(A, rowptr, colidx, vals, false);
    if (rc) { free_csr(A); return rc; }
    std::vector<int> ghosts, need(P, 0);
    rc = device_halo_plan(A->colidx, nnz, rows, global_rows, row_offset, ghosts);
    if (rc) { free_csr(A); return rc; }
    const int nghost = (int)ghosts.size();
    A->nghost = nghost;
    std::vector<int> ghost_local(nghost);
    ghost_owners(ghosts, global_rows, P, need.data(), ghost_local.data());
    A->recv_count = need;
    // need matrix: all[r * P + p] = rows rank r needs from rank p
    int *d_all = nullptr, *d_ids = nullptr;
    rc = rails::dev_alloc_t(&d_all, (size_t)P * P * sizeof(int));
    if (rc) { free_csr(A); return rc; }
    RAILS_CUDA_TRY(cudaMemcpyAsync(d_all + (size_t)c.rank * P, need.data(), P * sizeof(int),
                                   cudaMemcpyHostToDevice, c.stream));
    RAILS_NCCL_TRY(ncclAllGather(d_all + (size_t)c.rank * P, d_all, P, ncclInt, c.comm, c.stream));
    std::vector<int> all((size_t)P * P);
    RAILS_CUDA_TRY(cudaMemcpyAsync(all.data(), d_all, all.size() * sizeof(int), cudaMemcpyDeviceToHost, c.stream));
    RAILS_CUDA_TRY(cudaStreamSynchronize(c.stream));
    for (int p = 0; p < P; ++p)
    {
        A->send_count[p] = all[(size_t)p * P + c.rank];
        A->nsend += A->send_count[p];
    }
    rc = rails::dev_alloc_t(&d_ids, std::max<size_t>(ghost_local.size(), 1) * sizeof(int));
    if (rc) { free_csr(A); return rc; }
    rc = rails::dev_alloc_t(&A->send_rows, std::max<size_t>(A->nsend, 1) * sizeof(int));
    if (rc) { free_csr(A); return rc; }
    if (!ghost_local.empty())
        RAILS_CUDA_TRY(cudaMemcpyAsync(d_ids, ghost_local.data(), ghost_local.size() * sizeof(int),
                                       cudaMemcpyHostToDevice, c.stream));
    RAILS_NCCL_TRY(ncclGroupStart());
    {
        long long goff = 0, soff = 0;
        for (int p = 0; p < P; ++p)
        {
            if (need[p])
                RAILS_NCCL_TRY(ncclSend(d_ids + goff, need[p], ncclInt, p, c.comm, c.stream));
            if (A->send_count[p])
                RAILS_NCCL_TRY(ncclRecv(A->send_rows + soff, A->send_count[p], ncclInt, p, c.comm, c.stream));
            goff += need[p];
            soff += A->send_count[p];
        }
    }
    RAILS_NCCL_TRY(ncclGroupEnd());
    RAILS_CUDA_TRY(cudaStreamSynchronize(c.stream));
    rails::dev_free(d_all);
    rails::dev_free(d_ids);
    *out = A;
    return RAILS_OK;
}

extern "C" int rails_csr_free(rails_csr_t A)
{
    free_csr(A);
    return RAILS_OK;
}

// ---- element access and value updates -------------------------------------------

namespace
{

// position of entry (row, col) in the value array, or -1 (host-side lookup over the
// row's column indices; test-only path)
int locate_entry(rails_csr_s *A, long long row, long long col, long long *pos)
{
    if (A->partitioned)
        return rails::fail(RAILS_ERR_ARG, "csr element access is not available on a row-partitioned operator");
    if (row < 0 || row >= A->rows || col < 0 || col >= A->global_cols)
        return rails::fail(RAILS_ERR_ARG, "csr element (%lld, %lld) outside %lld x %lld", row, col, A->rows,
                           A->global_cols);
    int rp[2];
    RAILS_CUDA_TRY(cudaMemcpyAsync(rp, A->rowptr + row, 2 * sizeof(int), cudaMemcpyDeviceToHost, ctx().stream));
    RAILS_CUDA_TRY(cudaStreamSynchronize(ctx().stream));
    std::vector<int> cols(std::max(rp[1] - rp[0], 1));
    if (rp[1] > rp[0])
    {
        RAILS_CUDA_TRY(cudaMemcpyAsync(cols.data(), A->colidx + rp[0], (size_t)(rp[1] - rp[0]) * sizeof(int),
                                       cudaMemcpyDeviceToHost, ctx().stream));
        RAILS_CUDA_TRY(cudaStreamSynchronize(ctx().stream));
    }
    *pos = -1;
    for (int p = 0; p < rp[1] - rp[0]; ++p)
        if (cols[p] == (int)col) { *pos = rp[0] + p; break; }
    return RAILS_OK;
}

void drop_transpose(rails_csr_s *A)
{
    if (A->transposed) { free_csr(A->transposed); A->transposed = nullptr; }
}

__global__ void __launch_bounds__(256)
csr_to_dense_kernel(const int *__restrict__ rowptr, const int *__restrict__ colidx,
                    const double *__restrict__ vals, long long rows, double *__restrict__ out, long long ld)
{
    for (long long r = (long long)blockIdx.x * 256 + threadIdx.x; r < rows; r += (long long)gridDim.x * 256)
        for (int p = rowptr[r]; p < rowptr[r + 1]; ++p)
            out[r + (size_t)colidx[p] * ld] += vals[p];        // duplicates add up, as InsertGlobalValues sums them
}

}

// Test-only element access (the reference's Operator keeps `(m, n)` + operator()(i, j)
// for its typed tests: src/Epetra_OperatorWrapper.cpp:224-282): only entries that
// are part of the sparsity pattern can be read or written.
extern "C" int rails_csr_get_value(rails_csr_t A, long long row, long long col, double *value)
{
    RAILS_REQUIRE_READY();
    if (!A || !value) return rails::fail(RAILS_ERR_ARG, "csr_get_value: null argument");
    long long pos = -1;
    int rc = locate_entry(A, row, col, &pos);
    if (rc) return rc;
    *value = 0.0;
    if (pos < 0) return RAILS_OK;                   // outside the pattern: a structural zero
    RAILS_CUDA_TRY(cudaMemcpyAsync(value, A->vals + pos, sizeof(double), cudaMemcpyDeviceToHost, ctx().stream));
    RAILS_CUDA_TRY(cudaStreamSynchronize(ctx().stream));
    return RAILS_OK;
}

extern "C" int rails_csr_set_value(rails_csr_t A, long long row, long long col, double value)
{
    RAILS_REQUIRE_READY();
    if (!A) return rails::fail(RAILS_ERR_ARG, "csr_set_value: null operator");
    long long pos = -1;
    int rc = locate_entry(A, row, col, &pos);
    if (rc) return rc;
    if (pos < 0)
        return rails::fail(RAILS_ERR_ARG, "csr_set_value: (%lld, %lld) is not in the sparsity pattern", row, col);
    RAILS_CUDA_TRY(cudaMemcpyAsync(A->vals + pos, &value, sizeof(double), cudaMemcpyHostToDevice, ctx().stream));
    RAILS_CUDA_TRY(cudaStreamSynchronize(ctx().stream));
    drop_transpose(A);
    return RAILS_OK;
}

// New values on the same sparsity pattern, in the order of the arrays given to
// rails_csr_create (a continuation step A + delta I, or a re-assembled Jacobian):
// the row blocks, the renumbered columns and the halo plan are kept.
extern "C" int rails_csr_update_values(rails_csr_t A, const double *vals)
{
    RAILS_REQUIRE_READY();
    if (!A || (A->nnz > 0 && !vals)) return rails::fail(RAILS_ERR_ARG, "csr_update_values: null argument");
    if (A->nnz)
    {
        int rc = rails::copy_h2d(A->vals, vals, (size_t)A->nnz * sizeof(double));
        if (rc) return rc;
    }
    drop_transpose(A);
    return RAILS_OK;
}

// Out[:, oc0 .. oc0 + cols) = the operator as a dense block (small operators only:
// Operator::norm() of the reference forms A'A explicitly,
// src/Epetra_OperatorWrapper.cpp:115-145)
extern "C" int rails_csr_to_dense(rails_csr_t A, rails_mv_t Out, int oc0)
{
    RAILS_REQUIRE_READY();
    if (!A || !Out) return rails::fail(RAILS_ERR_ARG, "csr_to_dense: null argument");
    if (A->partitioned)
        return rails::fail(RAILS_ERR_ARG, "csr_to_dense is not available on a row-partitioned operator");
    if (Out->rows != A->rows || oc0 < 0 || oc0 + A->global_cols > Out->cols)
        return rails::fail(RAILS_ERR_ARG, "csr_to_dense: result must be %lld x %lld", A->rows, A->global_cols);
    int rc = rails_mv_fill(Out, oc0, (int)A->global_cols, 0.0);
    if (rc || A->rows == 0) return rc;
    const int nb = (int)std::min<long long>((A->rows + 255) / 256, (long long)ctx().sm_count * 8);
    RAILS_LAUNCH(csr_to_dense_kernel, nb, 256, 0, A->rowptr, A->colidx, A->vals, A->rows,
                 Out->data + (size_t)oc0 * Out->ld, Out->ld);
    RAILS_CHECK_LAUNCH();
    return RAILS_OK;
}

extern "C" long long rails_csr_rows(rails_csr_t A) { return A ? A->rows : -1; }
extern "C" long long rails_csr_nnz(rails_csr_t A) { return A ? A->nnz : -1; }

namespace
{

// A' of this rank's block as a CSR operator of its own, built on the device
// (transpose.cu).  For a row-partitioned A the block is rows x [owned | ghost] columns,
// so A' has ncols_local + nghost rows: the owned part of A'x comes out first, the
// ghost part is what the owners of those rows still have to add (reverse halo).
int build_transpose(rails_csr_s *A)
{
    if (A->transposed) return RAILS_OK;
    const long long m = A->rows, n = A->ncols_local + A->nghost, nnz = A->nnz;
    rails_csr_s *T = new rails_csr_s();
    T->rows = n; T->global_rows = n; T->global_cols = m; T->row_offset = 0; T->nnz = nnz;
    T->ncols_local = m;
    T->rowptr = T->colidx = T->blk_row = nullptr; T->vals = nullptr;
    T->send_rows = nullptr; T->send_buf = T->ghost_buf = nullptr;
    T->nghost = T->nsend = 0; T->halo_cols_cap = 0; T->transposed = nullptr; T->bytes = 0;
    T->partitioned = false;
    T->recv_count.assign(ctx().nranks, 0);
    T->send_count.assign(ctx().nranks, 0);
    int rc = rails::csr_transpose_device(m, n, nnz, A->rowptr, A->colidx, A->vals, &T->rowptr, &T->colidx, &T->vals);
    if (rc) { free_csr(T); return rc; }
    T->bytes = (size_t)(n + 1) * sizeof(int) + (size_t)nnz * (sizeof(int) + sizeof(double)) + 128;
    ctx().bytes_in_use.fetch_add((long long)T->bytes);
    // the block table is cut on the host from the row pointers (4 (n + 1) bytes back)
    std::vector<int> h_rowptr(n + 1);
    rc = rails::copy_d2h(h_rowptr.data(), T->rowptr, (size_t)(n + 1) * sizeof(int));
    if (!rc) rc = plan_blocks(T, h_rowptr.data());
    if (rc) { free_csr(T); return rc; }
    A->transposed = T;
    return RAILS_OK;
}

__global__ void __launch_bounds__(256)
add_rows_kernel(const int *__restrict__ rows, int n, const double *__restrict__ buf, long long ldb,
                double *__restrict__ Y, long long ldy)
{
    const int c = blockIdx.y;
    for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256)
        Y[rows[i] + (size_t)c * ldy] += buf[i + (size_t)c * ldb];
}

// Y = A' X on a row-partitioned A: Z = (local block)' X has ncols_local + nghost rows;
// its first ncols_local rows are this rank's own contribution to Y, the ghost rows go to
// their owners, who add what they receive peer by peer in rank order (a fixed order: the
// result does not depend on arrival times).  The plan is the forward halo plan with the
// roles of the lists exchanged (Epetra: Export with Add over the column map).
int transposed_apply_partitioned(rails_csr_s *A, const double *X, long long ldx, int ncols, double *Y, long long ldy)
{
    rails::Context &c = ctx();
    rails_csr_s *T = A->transposed;
    const long long nz = T->rows;                       // ncols_local + nghost
    double *Z = nullptr;
    int rc = rails::dev_alloc_t(&Z, (size_t)std::max<long long>(nz, 1) * ncols * sizeof(double));
    if (rc) return rc;
    rc = launch_spmm(T, X, ldx, ncols, Z, nz);
    if (!rc) rc = ensure_halo_cols(A, ncols);
    if (rc) { rails::dev_free(Z); return rc; }
    if (A->ncols_local)
        RAILS_CUDA_TRY(cudaMemcpy2DAsync(Y, (size_t)ldy * sizeof(double), Z, (size_t)nz * sizeof(double),
                                         (size_t)A->ncols_local * sizeof(double), ncols, cudaMemcpyDeviceToDevice,
                                         c.stream));
    RAILS_NCCL_TRY(ncclGroupStart());
    long long soff = 0, goff = 0;
    for (int p = 0; p < c.nranks; ++p)
    {
        for (int col = 0; col < ncols; ++col)
        {
            if (A->recv_count[p])           // ghost rows of Z go back to their owner
                RAILS_NCCL_TRY(ncclSend(Z + A->ncols_local + goff + (size_t)col * nz, A->recv_count[p], ncclDouble, p,
                                        c.comm, c.stream));
            if (A->send_count[p])           // contributions to the rows this rank sends in the forward halo
                RAILS_NCCL_TRY(ncclRecv(A->send_buf + soff + (size_t)col * A->nsend, A->send_count[p], ncclDouble, p,
                                        c.comm, c.stream));
        }
        soff += A->send_count[p];
        goff += A->recv_count[p];
    }
    RAILS_NCCL_TRY(ncclGroupEnd());
    soff = 0;
    for (int p = 0; p < c.nranks; ++p)
    {
        if (A->send_count[p])
        {
            dim3 grid(std::min((A->send_count[p] + 255) / 256, c.sm_count * 4), ncols);
            RAILS_LAUNCH(add_rows_kernel, grid, 256, 0, (const int *)(A->send_rows + soff), A->send_count[p],
                         (const double *)(A->send_buf + soff), (long long)A->nsend, Y, ldy);
        }
        soff += A->send_count[p];
    }
    RAILS_CHECK_LAUNCH();
    rails::dev_free(Z);          // stream-ordered reuse: later work on this stream comes after the sends
    return RAILS_OK;
}

}

extern "C" int rails_csr_spmm(rails_csr_t A, int trans, rails_mv_t X, int xc0, int ncols,
                              rails_mv_t Y, int yc0)
{
    RAILS_REQUIRE_READY();
    if (!A || !X || !Y)
        return rails::fail(RAILS_ERR_ARG, "spmm: null handle");
    if (xc0 < 0 || ncols < 0 || xc0 + ncols > X->cols || yc0 < 0 || yc0 + ncols > Y->cols)
        return rails::fail(RAILS_ERR_ARG, "spmm: column range outside capacity");
    if (X->data == Y->data)
        return rails::fail(RAILS_ERR_ARG, "spmm: X and Y must be different buffers");
    rails_csr_s *Op = A;
    if (trans)
    {
        int rc = build_transpose(A);
        if (rc) return rc;
        Op = A->transposed;
        if (A->partitioned)
        {
            if (X->rows != A->rows || Y->rows != A->ncols_local)
                return rails::fail(RAILS_ERR_ARG, "spmm: transposed operator %lld x %lld applied to %lld rows -> %lld rows",
                                   A->ncols_local, A->rows, X->rows, Y->rows);
            if (ncols == 0) return RAILS_OK;
            return transposed_apply_partitioned(A, X->data + (size_t)xc0 * X->ld, X->ld, ncols,
                                                Y->data + (size_t)yc0 * Y->ld, Y->ld);
        }
    }
    if (X->rows != Op->ncols_local || Y->rows != Op->rows)
        return rails::fail(RAILS_ERR_ARG, "spmm: operator %lld x %lld applied to %lld rows -> %lld rows",
                           Op->rows, Op->ncols_local, X->rows, Y->rows);
    const double *Xd = X->data + (size_t)xc0 * X->ld;
    if (Op->nghost || Op->nsend)
    {
        int rc = halo_exchange(Op, Xd, X->ld, ncols);
        if (rc) return rc;
    }
    return launch_spmm(Op, Xd, X->ld, ncols, Y->data + (size_t)yc0 * Y->ld, Y->ld);
}

// sqrt(lambda_max(A'A)) by power iteration (EpetraOp::norm forms A'A and
// asks an eigensolver for the largest eigenvalue, :115-145).
extern "C" int rails_csr_norm2(rails_csr_t A, double *out)
{
    RAILS_REQUIRE_READY();
    if (!A || !out)
        return rails::fail(RAILS_ERR_ARG, "csr_norm2: null argument");
    int rc = build_transpose(A);
    if (rc) return rc;
    // on a row-partitioned operator x and y are distributed like its rows (the dots
    // below then reduce over the ranks) and the start vector is the same global vector
    // whatever the partition
    const int flags = A->partitioned ? 0 : RAILS_MV_REPLICATED;
    rails_mv_t x = nullptr, y = nullptr;
    rc = rails_mv_alloc(A->ncols_local, 1, flags, &x);
    if (rc) return rc;
    rc = rails_mv_alloc(A->rows, 1, flags, &y);
    if (rc) { rails_mv_free(x); return rc; }
    rails_mv_random(x, 0, 1, A->partitioned ? A->row_offset : 0, 12345u);
    double lambda = 0.0, prev = -1.0;
    for (int it = 0; it < 1000; ++it)
    {
        double nx = 0.0;
        rc = rails_mv_dot(x, 0, 1, x, 0, 1, &nx, 1);
        if (rc) break;
        if (nx == 0.0) { lambda = 0.0; break; }
        rails_mv_scal(x, 0, 1, 1.0 / sqrt(nx));
        rc = rails_csr_spmm(A, 0, x, 0, 1, y, 0);
        if (rc) break;
        rc = rails_mv_dot(y, 0, 1, y, 0, 1, &lambda, 1);   // |A x|^2 = x'A'A x
        if (rc) break;
        rc = rails_csr_spmm(A, 1, y, 0, 1, x, 0);
        if (rc) break;
        if (fabs(lambda - prev) <= 1e-14 * fabs(lambda)) break;
        prev = lambda;
    }
    rails_mv_free(x);
    rails_mv_free(y);
    if (rc) return rc;
    *out = sqrt(lambda);
    return RAILS_OK;
}
