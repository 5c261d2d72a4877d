// Internal declarations shared by the .cu files of librails_cuda.so.
#ifndef RAILS_B200_INTERNAL_H
#define RAILS_B200_INTERNAL_H

#include <cuda_runtime.h>
#include <nccl.h>

#include <atomic>
#include <cstdio>
#include <string>
#include <vector>

#include "rails_cuda.h"

struct rails_mv_s
{
    double *data;
    long long rows;   // local rows
    long long ld;     // leading dimension, multiple of 16 doubles
    int cols;         // capacity in columns
    int flags;        // RAILS_MV_DISTRIBUTED / RAILS_MV_REPLICATED
    size_t bytes;
};

struct rails_csr_s
{
    long long rows;          // local rows
    long long global_rows;
    long long global_cols;
    long long row_offset;
    long long nnz;
    long long ncols_local;   // owned columns (= local rows of X)
    // device arrays; column indices renumbered [owned | ghost]
    int *rowptr;
    int *colidx;
    double *vals;
    // row blocks for the staged kernel: block b covers rows
    // [blk_row[b], blk_row[b+1])
    int *blk_row;
    int nblk;
    int max_blk_nnz;
    int blk_cap;                         // staging capacity (nnz) the blocks were cut for
    // halo plan (distributed only)
    int nghost;                          // ghost rows of X needed here
    std::vector<int> recv_count;         // per peer, ghost rows received
    std::vector<int> send_count;         // per peer, owned rows sent
    int *send_rows;                      // device: owned local row ids to pack
    int nsend;
    double *send_buf;                    // device staging, nsend x ncols
    double *ghost_buf;                   // device, nghost x ncols
    int halo_cols_cap;                   // columns the staging buffers hold
    bool partitioned;                    // rows != global rows (no transpose available)
    rails_csr_s *transposed;             // built lazily from the device arrays
    size_t bytes;
};

namespace rails
{

// Peer-memory mailboxes for small reductions over NVLink/NVSwitch: every rank
// maps every other rank's mailbox (CUDA IPC) and the reduction kernel stores
// its contribution straight into the peers' memory (exchange.cu).
constexpr int MB_MAX_RANKS = 8;
constexpr int MB_SLOTS = 4;
constexpr int MB_MAXN = 2048;          // doubles per contribution
struct Mailbox
{
    double *data[MB_MAX_RANKS];               // rank p's data area: [slot][writer][MB_MAXN]
    unsigned long long *flags[MB_MAX_RANKS];  // rank p's flags:     [slot][writer]
    int rank, nranks;
};

struct Context
{
    bool ready = false;
    int device = -1;
    int sm_count = 148;
    cudaStream_t stream = nullptr;
    cudaStream_t comm_stream = nullptr;
    // reduction workspace (device), small result staging (device + pinned host)
    double *d_partial = nullptr;
    size_t partial_cap = 0;       // doubles
    double *d_small = nullptr;
    size_t small_cap = 0;         // doubles
    double *h_small = nullptr;    // pinned
    double *d_coef = nullptr;     // uploaded coefficient blocks for gemm
    size_t coef_cap = 0;
    size_t coef_pos = 0;
    double *h_ring = nullptr;     // pinned staging ring for small H2D operands
    size_t ring_cap = 0;
    size_t ring_pos = 0;
    void *h_stage[2] = {nullptr, nullptr};   // pinned chunks for large pageable transfers
    cudaEvent_t stage_ev[2] = {nullptr, nullptr};
    // communicator
    ncclComm_t comm = nullptr;
    int rank = 0;
    int nranks = 1;
    // peer mailboxes (nranks > 1 and all peers reachable over NVLink)
    bool p2p = false;
    Mailbox mb;
    void *mb_local = nullptr;
    void *mb_mapped[MB_MAX_RANKS] = {nullptr};
    unsigned long long mb_seq = 0;           // exchanges issued so far (same on every rank)
    // raised by a kernel whose wait for a peer timed out: host-mapped pinned
    // memory, so that every host read-back can test it without a copy
    int *h_mb_error = nullptr;
    int *d_mb_error = nullptr;               // device address of *h_mb_error
    long long mb_timeout_cycles = 0;         // RAILS_MAILBOX_TIMEOUT_S (default 30 s) in SM clocks
    // rails_mv_lanczos_begin -> rails_mv_lanczos_run hand-over: scratch that
    // survives other library calls, and what the pending first projection was for
    double *d_lz = nullptr;
    size_t lz_cap = 0;
    bool lz_pending = false;
    const double *lz_q = nullptr;
    int lz_total = 0;
    std::atomic<long long> launches{0};
    std::atomic<long long> host_syncs{0};      // stream synchronisations the library made (host waits for the GPU)
    std::atomic<long long> errors{0};        // calls that failed since the library was loaded
    std::atomic<long long> bytes_in_use{0};
    std::vector<cudaEvent_t> events;
    std::string last_error;
};

Context &ctx();
int fail(int code, const char *fmt, ...);
// Device blocks from a size-keyed cache over cudaMalloc (context.cu): freed
// blocks are kept and handed out again to requests of (nearly) the same size.
// Everything using them is ordered on the one library stream, which makes
// immediate reuse safe.
int dev_alloc(void **p, size_t bytes);
template<class T> int dev_alloc_t(T **p, size_t bytes)
{
    void *v = nullptr;
    int rc = dev_alloc(&v, bytes);
    *p = static_cast<T *>(v);
    return rc;
}
void dev_free(void *p);
void dev_cache_release();
int ensure_partial(size_t doubles);
int ensure_small(size_t doubles);
// Copies a small host block (column-major k x c, ldc) into the pinned ring and
// enqueues its upload; *d_out is a compact (ld = k) device copy that stays
// valid until the ring wraps (thousands of calls later, guarded by a sync).
int upload_small(const double *host, int k, int c, int ldc, const double **d_out);
// Large transfers between PAGEABLE host memory and the device: chunks are
// moved through two pinned buffers, the host-side memcpy of a chunk is split
// over a few threads and overlaps the DMA of the previous chunk.  Synchronous
// (the host block may be reused / is complete on return).
int copy_h2d(void *dst, const void *src, size_t bytes);
int copy_d2h(void *dst, const void *src, size_t bytes);
int copy_d2h_2d(void *dst, size_t dst_ld, const void *src, size_t src_ld, size_t col_bytes, int ncols);
// transpose.cu: CSR arrays of A' built on the device (outputs from dev_alloc)
int csr_transpose_device(long long rows, long long ncols, long long nnz, const int *rowptr, const int *colidx,
                         const double *vals, int **rowptr_t, int **colidx_t, double **vals_t);

// Finishes a reduction and, with a communicator, sums it over the ranks, in ONE
// kernel: out[i] = sum over ranks of (sum_b partial[b * n + i]), b < nblocks.
// `partial` may alias `out` when nblocks == 1 (plain all-reduce of n doubles).
// With T != NULL (k x k, compact) it continues with the Lanczos coefficient
// products: coef = [T out[k:2k) ; T out[0:k) ; out[2k:n)].
// Uses the peer mailboxes when available, otherwise a second-stage kernel + NCCL.
// Scaling options (device-resident recurrences): with norm_out != NULL the LAST of
// the n values is a squared norm: it is stored to *norm_out and the other n - 1
// values are multiplied by 1 / sqrt(|it|) before they are stored / used for coef;
// with scale_sq != NULL every value is multiplied by 1 / sqrt(|*scale_sq|).
// With coef and alpha_out != NULL, *alpha_out = out . coef (= q'Rq of the Lanczos step).
int finish_reduction(const double *partial, int nblocks, int n, double *out, bool over_ranks,
                     const double *T = nullptr, int k = 0, double *coef = nullptr,
                     double *norm_out = nullptr, const double *scale_sq = nullptr,
                     double *alpha_out = nullptr);
// coef = [T g[k:2k) ; T g[0:k) ; g[2k:n)] on the device (T compact k x k)
int lanczos_coef(const double *T, int k, int n, const double *g, double *coef, double *alpha_out = nullptr);
// After a stream synchronisation that returned reduction results to the host:
// non-zero (flag cleared, error count bumped so that BackendHealth stops the solve)
// when a peer did not answer an exchange in time -- the values just read are NaN.
int check_exchange_error(const char *where);
int setup_mailboxes();
void release_mailboxes();

// device-side kernels callable across translation units
int dot_device(const double *X, long long ldx, int a, const double *Y,
               long long ldy, int b, long long rows, double *d_out, int ldo,
               bool reduce_over_ranks);
// X' Y with both a and b above 8 on the FP64 tensor cores (gram.cu); operands
// 16-byte aligned with even leading dimensions
int gram_device(const double *X, long long ldx, int a, const double *Y, long long ldy, int b,
                long long rows, double *d_out, int ldo);
// Out = alpha V C + beta Out with more than 8 result columns on the FP64 tensor
// cores (gram.cu); operands 16-byte aligned with even leading dimensions
bool fp64_tensor_enabled();
bool gemm_dmma_fits(int k, int c);
int gemm_dmma_device(const double *V, long long ldv, int k, const double *dC, int ldc, int c,
                     double alpha, double beta, double *Out, long long ldo, long long rows);
int gemm_device(const double *V, long long ldv, int k, const double *dC, int ldc,
                int c, double alpha, double beta, double *Out, long long ldo,
                long long rows);

}

#define RAILS_CUDA_TRY(expr)                                                   \
    do {                                                                       \
        cudaError_t e__ = (expr);                                              \
        if (e__ != cudaSuccess)                                                \
            return rails::fail(RAILS_ERR_CUDA, "%s:%d %s -> %s", __FILE__,     \
                               __LINE__, #expr, cudaGetErrorString(e__));      \
    } while (0)

#define RAILS_NCCL_TRY(expr)                                                   \
    do {                                                                       \
        ncclResult_t r__ = (expr);                                             \
        if (r__ != ncclSuccess)                                                \
            return rails::fail(RAILS_ERR_CUDA, "%s:%d %s -> %s", __FILE__,     \
                               __LINE__, #expr, ncclGetErrorString(r__));      \
    } while (0)

#define RAILS_REQUIRE_READY()                                                  \
    do {                                                                       \
        if (!rails::ctx().ready)                                               \
            return rails::fail(RAILS_ERR_STATE,                                \
                               "rails_cuda_init has not been called");         \
    } while (0)

// every kernel launch goes through this so gpu_launches is a real count
#define RAILS_LAUNCH(kernel, grid, block, smem, ...)                           \
    do {                                                                       \
        kernel<<<(grid), (block), (smem), rails::ctx().stream>>>(__VA_ARGS__); \
        rails::ctx().launches.fetch_add(1, std::memory_order_relaxed);         \
    } while (0)

#define RAILS_CHECK_LAUNCH() RAILS_CUDA_TRY(cudaGetLastError())

// every cudaStreamSynchronize of the library is counted (rails_cuda_host_syncs):
// the number of times per iteration the host waits for the device is a reported figure
namespace rails
{
inline cudaError_t stream_sync_counted(cudaStream_t s)
{
    ctx().host_syncs.fetch_add(1, std::memory_order_relaxed);
    return (cudaStreamSynchronize)(s);
}
}
#define cudaStreamSynchronize(s) rails::stream_sync_counted(s)

#endif
