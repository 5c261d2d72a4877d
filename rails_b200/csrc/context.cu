// Context, communicator and multivector storage of librails_cuda.so.
#include "internal.h"

#include <sched.h>
#include <cctype>
#include <cstdio>
#include <string>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <map>
#include <thread>
#include <unordered_map>

#include <cstdarg>
#include <cstring>

namespace rails
{

Context &ctx()
{
    static Context c;
    return c;
}

int fail(int code, const char *fmt, ...)
{
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    ctx().last_error = buf;
    ctx().errors.fetch_add(1);
    fprintf(stderr, "rails_cuda error: %s\n", buf);
    return code;
}

static int grow(double **p, size_t *cap, size_t need)
{
    if (need <= *cap)
        return RAILS_OK;
    Context &c = ctx();
    size_t ncap = *cap ? *cap : 4096;
    while (ncap < need)
        ncap *= 2;
    if (*p)
    {
        // stream-ordered: earlier kernels that still use the old block finish first
        RAILS_CUDA_TRY(cudaStreamSynchronize(c.stream));
        RAILS_CUDA_TRY(cudaFree(*p));
    }
    RAILS_CUDA_TRY(cudaMalloc(p, ncap * sizeof(double)));
    *cap = ncap;
    return RAILS_OK;
}

int ensure_partial(size_t doubles) { return grow(&ctx().d_partial, &ctx().partial_cap, doubles); }

int check_exchange_error(const char *where)
{
    Context &c = ctx();
    if (!c.h_mb_error || !*(volatile int *)c.h_mb_error)
        return RAILS_OK;
    *(volatile int *)c.h_mb_error = 0;
    return fail(RAILS_ERR_CUDA, "%s: a peer did not answer a reduction within the mailbox timeout "
                "(RAILS_MAILBOX_TIMEOUT_S); the values returned are NaN", where);
}

// ---- device block cache --------------------------------------------------------
// Multivector buffers (V, AV, MV at capacity; the m x c temporaries the wrapper
// API creates every iteration) and operator arrays come from cudaMalloc and go
// back into a size-keyed free list instead of to the driver.  Round 1 used the
// stream-ordered pool (cudaMallocAsync): identical solves then differed by single
// iterations of 100-850 ms (restart temporaries of a few GB made the pool remap
// physical memory; scripts/window_variability.py, profiles/r02_window_variability.md).
// Everything that touches these buffers is enqueued on the ONE library stream, so
// a block can be handed out again as soon as it has been returned.
namespace
{
std::multimap<size_t, void *> g_free_blocks;        // size -> block
std::unordered_map<void *, size_t> g_live_blocks;   // block handed out -> its size
size_t g_free_bytes = 0;

size_t round_block(size_t bytes)
{
    // 512 B granules below 1 MiB, 2 MiB granules above (the driver's own page size)
    const size_t g = bytes < (1u << 20) ? 512 : (2u << 20);
    return (std::max<size_t>(bytes, 1) + g - 1) / g * g;
}
}

void dev_cache_release()
{
    Context &c = ctx();
    if (c.stream) cudaStreamSynchronize(c.stream);
    for (auto &kv : g_free_blocks)
        cudaFree(kv.second);
    g_free_blocks.clear();
    g_free_bytes = 0;
}

int dev_alloc(void **p, size_t bytes)
{
    const size_t want = round_block(bytes);
    // smallest cached block that fits without wasting more than a quarter of it
    auto it = g_free_blocks.lower_bound(want);
    if (it != g_free_blocks.end() && it->first <= want + want / 4 + (1u << 20))
    {
        *p = it->second;
        g_live_blocks[*p] = it->first;
        g_free_bytes -= it->first;
        g_free_blocks.erase(it);
        return RAILS_OK;
    }
    cudaError_t e = cudaMalloc(p, want);
    if (e == cudaErrorMemoryAllocation)
    {
        // cached blocks of other sizes (a capacity growth leaves the old V, AV
        // behind) are in the way: give them back and retry
        cudaGetLastError();
        dev_cache_release();
        e = cudaMalloc(p, want);
    }
    if (e != cudaSuccess)
    {
        cudaGetLastError();
        return fail(RAILS_ERR_CUDA, "cudaMalloc(%zu bytes): %s", want, cudaGetErrorString(e));
    }
    g_live_blocks[*p] = want;
    return RAILS_OK;
}

void dev_free(void *p)
{
    if (!p) return;
    auto live = g_live_blocks.find(p);
    if (live == g_live_blocks.end())
    {
        fail(RAILS_ERR_STATE, "dev_free: %p was not handed out by dev_alloc", p);
        return;
    }
    const size_t granted = live->second;
    g_live_blocks.erase(live);
    g_free_blocks.emplace(granted, p);
    g_free_bytes += granted;
    // bound what the cache may hoard: beyond half of the device memory the
    // largest blocks go back to the driver (they are the least likely to be reused)
    static size_t limit = 0;
    if (!limit)
    {
        size_t fr = 0, tot = 0;
        limit = cudaMemGetInfo(&fr, &tot) == cudaSuccess ? tot / 2 : (size_t)64 << 30;
    }
    while (g_free_bytes > limit && !g_free_blocks.empty())
    {
        auto last = std::prev(g_free_blocks.end());
        cudaStreamSynchronize(ctx().stream);
        cudaFree(last->second);
        g_free_bytes -= last->first;
        g_free_blocks.erase(last);
    }
}

int upload_small(const double *host, int k, int c, int ldc, const double **d_out)
{
    Context &cx = ctx();
    const size_t n = (size_t)k * c;
    if (n > cx.ring_cap)
    {
        // oversized block: grow both rings (rare: warm-start / restart blocks)
        RAILS_CUDA_TRY(cudaStreamSynchronize(cx.stream));
        size_t ncap = cx.ring_cap ? cx.ring_cap : (1u << 20);
        while (ncap < n) ncap *= 2;
        if (cx.h_ring) RAILS_CUDA_TRY(cudaFreeHost(cx.h_ring));
        if (cx.d_coef) RAILS_CUDA_TRY(cudaFree(cx.d_coef));
        RAILS_CUDA_TRY(cudaMallocHost(&cx.h_ring, ncap * sizeof(double)));
        RAILS_CUDA_TRY(cudaMalloc(&cx.d_coef, ncap * sizeof(double)));
        cx.ring_cap = cx.coef_cap = ncap;
        cx.ring_pos = 0;
    }
    if (cx.ring_pos + n > cx.ring_cap)
    {
        // wrap: everything staged so far must have been consumed
        RAILS_CUDA_TRY(cudaStreamSynchronize(cx.stream));
        cx.ring_pos = 0;
    }
    double *h = cx.h_ring + cx.ring_pos;
    double *d = cx.d_coef + cx.ring_pos;
    for (int j = 0; j < c; ++j)
        memcpy(h + (size_t)j * k, host + (size_t)j * ldc, (size_t)k * sizeof(double));
    RAILS_CUDA_TRY(cudaMemcpyAsync(d, h, n * sizeof(double), cudaMemcpyHostToDevice, cx.stream));
    cx.ring_pos += (n + 15) / 16 * 16;
    *d_out = d;
    return RAILS_OK;
}

namespace
{
constexpr size_t STAGE_CHUNK = 32u << 20;
constexpr int STAGE_THREADS = 6;

int ensure_stage()
{
    Context &c = ctx();
    for (int i = 0; i < 2; ++i)
        if (!c.h_stage[i])
        {
            RAILS_CUDA_TRY(cudaMallocHost(&c.h_stage[i], STAGE_CHUNK));
            RAILS_CUDA_TRY(cudaEventCreateWithFlags(&c.stage_ev[i], cudaEventDisableTiming));
        }
    return RAILS_OK;
}

// memcpy split over a few threads (one core moves ~10 GB/s, PCIe 5 x16 ~50)
void parallel_memcpy(void *dst, const void *src, size_t bytes)
{
    if (bytes < (4u << 20))
    {
        memcpy(dst, src, bytes);
        return;
    }
    std::thread th[STAGE_THREADS];
    const size_t part = (bytes / STAGE_THREADS + 4095) & ~(size_t)4095;
    int n = 0;
    for (size_t off = 0; off < bytes; off += part, ++n)
    {
        const size_t len = std::min(part, bytes - off);
        th[n] = std::thread([=] { memcpy((char *)dst + off, (const char *)src + off, len); });
    }
    for (int i = 0; i < n; ++i) th[i].join();
}
}

int copy_h2d(void *dst, const void *src, size_t bytes)
{
    Context &c = ctx();
    if (bytes < (8u << 20))
    {
        RAILS_CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, c.stream));
        RAILS_CUDA_TRY(cudaStreamSynchronize(c.stream));
        return RAILS_OK;
    }
    int rc = ensure_stage();
    if (rc) return rc;
    int b = 0;
    for (size_t off = 0; off < bytes; off += STAGE_CHUNK, b ^= 1)
    {
        const size_t len = std::min(STAGE_CHUNK, bytes - off);
        RAILS_CUDA_TRY(cudaEventSynchronize(c.stage_ev[b]));       // buffer b is free again
        parallel_memcpy(c.h_stage[b], (const char *)src + off, len);
        RAILS_CUDA_TRY(cudaMemcpyAsync((char *)dst + off, c.h_stage[b], len, cudaMemcpyHostToDevice, c.stream));
        RAILS_CUDA_TRY(cudaEventRecord(c.stage_ev[b], c.stream));
    }
    RAILS_CUDA_TRY(cudaStreamSynchronize(c.stream));
    return RAILS_OK;
}

// `ncols` columns of `col_bytes` each (strides dst_ld / src_ld bytes) through the two
// pinned buffers as ONE pipeline: piece i+1 is in flight while piece i is copied out of
// its buffer, across column boundaries too.
int copy_d2h_2d(void *dst, size_t dst_ld, const void *src, size_t src_ld, size_t col_bytes, int ncols)
{
    Context &c = ctx();
    if (col_bytes == 0 || ncols <= 0)
        return RAILS_OK;
    if (dst_ld == col_bytes && src_ld == col_bytes)
    {
        // contiguous on both sides: one long column
        col_bytes *= (size_t)ncols;
        dst_ld = src_ld = col_bytes;
        ncols = 1;
    }
    if (col_bytes * (size_t)ncols < (8u << 20))
    {
        RAILS_CUDA_TRY(cudaMemcpy2DAsync(dst, dst_ld, src, src_ld, col_bytes, ncols, cudaMemcpyDeviceToHost, c.stream));
        RAILS_CUDA_TRY(cudaStreamSynchronize(c.stream));
        return RAILS_OK;
    }
    int rc = ensure_stage();
    if (rc) return rc;
    // pieces: a slice of one tall column, or several whole short columns packed densely
    struct Piece { char *dst; size_t len; int cols; };
    Piece prev = {nullptr, 0, 0};
    auto drain = [&](int buf) -> int {
        RAILS_CUDA_TRY(cudaEventSynchronize(c.stage_ev[buf]));
        if (prev.cols <= 1)
            parallel_memcpy(prev.dst, c.h_stage[buf], prev.len);
        else
            for (int q = 0; q < prev.cols; ++q)
                memcpy(prev.dst + (size_t)q * dst_ld, (char *)c.h_stage[buf] + (size_t)q * col_bytes, col_bytes);
        return RAILS_OK;
    };
    int b = 0;
    const int pack = col_bytes <= STAGE_CHUNK / 2 ? (int)(STAGE_CHUNK / col_bytes) : 1;
    for (int j = 0; j < ncols; j += pack)
    {
        const int nc = std::min(pack, ncols - j);
        for (size_t off = 0; off < col_bytes; off += STAGE_CHUNK, b ^= 1)
        {
            const size_t len = std::min(STAGE_CHUNK, col_bytes - off);
            if (nc > 1)
                RAILS_CUDA_TRY(cudaMemcpy2DAsync(c.h_stage[b], col_bytes, (const char *)src + (size_t)j * src_ld, src_ld,
                                                 col_bytes, nc, cudaMemcpyDeviceToHost, c.stream));
            else
                RAILS_CUDA_TRY(cudaMemcpyAsync(c.h_stage[b], (const char *)src + (size_t)j * src_ld + off, len,
                                               cudaMemcpyDeviceToHost, c.stream));
            RAILS_CUDA_TRY(cudaEventRecord(c.stage_ev[b], c.stream));
            if (prev.dst)
            {
                rc = drain(b ^ 1);
                if (rc) return rc;
            }
            prev.dst = (char *)dst + (size_t)j * dst_ld + off;
            prev.len = len;
            prev.cols = nc;
        }
    }
    if (prev.dst)
        return drain(b ^ 1);
    return RAILS_OK;
}

int copy_d2h(void *dst, const void *src, size_t bytes)
{
    return copy_d2h_2d(dst, bytes, src, bytes, bytes, 1);
}

int ensure_small(size_t doubles)
{
    Context &c = ctx();
    if (doubles <= c.small_cap)
        return RAILS_OK;
    size_t ncap = c.small_cap ? c.small_cap : 65536;
    while (ncap < doubles)
        ncap *= 2;
    RAILS_CUDA_TRY(cudaStreamSynchronize(c.stream));
    if (c.d_small) RAILS_CUDA_TRY(cudaFree(c.d_small));
    if (c.h_small) RAILS_CUDA_TRY(cudaFreeHost(c.h_small));
    RAILS_CUDA_TRY(cudaMalloc(&c.d_small, ncap * sizeof(double)));
    RAILS_CUDA_TRY(cudaMallocHost(&c.h_small, ncap * sizeof(double)));
    c.small_cap = ncap;
    return RAILS_OK;
}

}

using rails::ctx;

namespace
{
// RAILS_NUMA_BIND=1: restrict the calling thread (and the threads it creates later: the
// staging copies) to the CPUs next to the GPU, as listed by the kernel in
// /sys/bus/pci/devices/<bus id>/local_cpulist, so that pinned staging buffers and the
// caller's arrays are first-touched on the GPU's NUMA node.  Off by default: measured
// on the 8-GPU box of this pool (profiles/r02_scaling.md).  Failures are ignored.
void bind_to_gpu_numa_node(int device)
{
    const char *e = getenv("RAILS_NUMA_BIND");
    if (!(e && e[0] == '1'))
        return;
    char bus[32] = {0};
    if (cudaDeviceGetPCIBusId(bus, sizeof(bus), device) != cudaSuccess)
    {
        cudaGetLastError();
        return;
    }
    for (char *p = bus; *p; ++p) *p = (char)tolower(*p);
    const std::string path = std::string("/sys/bus/pci/devices/") + bus + "/local_cpulist";
    FILE *f = fopen(path.c_str(), "r");
    if (!f) return;
    char line[4096] = {0};
    const bool got = fgets(line, sizeof(line), f) != nullptr;
    fclose(f);
    if (!got) return;
    cpu_set_t allowed, want;
    if (sched_getaffinity(0, sizeof(allowed), &allowed) != 0) return;
    CPU_ZERO(&want);
    int n = 0;
    for (char *tok = strtok(line, ",\n"); tok; tok = strtok(nullptr, ",\n"))
    {
        int a = 0, b = 0;
        const int parts = sscanf(tok, "%d-%d", &a, &b);
        if (parts == 1) b = a;
        if (parts < 1) continue;
        for (int cpu = a; cpu <= b && cpu < CPU_SETSIZE; ++cpu)
            if (CPU_ISSET(cpu, &allowed)) { CPU_SET(cpu, &want); ++n; }
    }
    if (n >= 2)
        sched_setaffinity(0, sizeof(want), &want);
    if (getenv("RAILS_NUMA_VERBOSE"))
        fprintf(stderr, "rails_cuda_init: device %d (%s) local cpus %d of %d allowed\n", device, bus, n,
                CPU_COUNT(&allowed));
}
}

extern "C" int rails_cuda_init(int device)
{
    rails::Context &c = ctx();
    if (c.ready)
        return RAILS_OK;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return rails::fail(RAILS_ERR_CUDA,
                           "no CUDA device available (%s); this backend has no CPU fallback",
                           cudaGetErrorString(e));
    if (device < 0 || device >= ndev)
        return rails::fail(RAILS_ERR_ARG, "device %d out of range (0..%d)", device, ndev - 1);
    RAILS_CUDA_TRY(cudaSetDevice(device));
    c.device = device;
    bind_to_gpu_numa_node(device);
    cudaDeviceProp prop;
    RAILS_CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    c.sm_count = prop.multiProcessorCount;
    RAILS_CUDA_TRY(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
    RAILS_CUDA_TRY(cudaStreamCreateWithFlags(&c.comm_stream, cudaStreamNonBlocking));
    c.events.assign(64, nullptr);
    RAILS_CUDA_TRY(cudaHostAlloc((void **)&c.h_mb_error, sizeof(int), cudaHostAllocMapped));
    *c.h_mb_error = 0;
    RAILS_CUDA_TRY(cudaHostGetDevicePointer((void **)&c.d_mb_error, c.h_mb_error, 0));
    {
        const char *e = getenv("RAILS_MAILBOX_TIMEOUT_S");
        double seconds = e ? atof(e) : 30.0;
        if (!(seconds > 0.0)) seconds = 30.0;
        c.mb_timeout_cycles = (long long)(seconds * 1e3 * (double)prop.clockRate);   // clockRate is in kHz
    }
    c.ready = true;
    int rc = rails::ensure_small(65536);
    if (rc) return rc;
    rc = rails::ensure_partial(1 << 20);
    if (rc) return rc;
    {
        const double dummy = 0.0;
        const double *d = nullptr;
        rc = rails::upload_small(&dummy, 1, 1, 1, &d);   // allocates the rings
        c.ring_pos = 0;
    }
    return rc;
}

extern "C" int rails_cuda_finalize(void)
{
    rails::Context &c = ctx();
    if (!c.ready)
        return RAILS_OK;
    cudaStreamSynchronize(c.stream);
    rails::dev_cache_release();
    rails::release_mailboxes();
    if (c.comm) { ncclCommDestroy(c.comm); c.comm = nullptr; }
    for (auto &ev : c.events)
        if (ev) { cudaEventDestroy(ev); ev = nullptr; }
    if (c.d_partial) cudaFree(c.d_partial);
    if (c.d_small) cudaFree(c.d_small);
    if (c.h_small) cudaFreeHost(c.h_small);
    if (c.d_coef) cudaFree(c.d_coef);
    if (c.h_ring) cudaFreeHost(c.h_ring);
    if (c.d_lz) cudaFree(c.d_lz);
    c.d_lz = nullptr; c.lz_cap = 0; c.lz_pending = false;
    if (c.h_mb_error) cudaFreeHost(c.h_mb_error);
    c.h_mb_error = c.d_mb_error = nullptr;
    for (int i = 0; i < 2; ++i)
    {
        if (c.h_stage[i]) cudaFreeHost(c.h_stage[i]);
        if (c.stage_ev[i]) cudaEventDestroy(c.stage_ev[i]);
        c.h_stage[i] = nullptr; c.stage_ev[i] = nullptr;
    }
    c.h_ring = nullptr; c.ring_cap = c.ring_pos = 0;
    c.d_partial = c.d_small = c.h_small = c.d_coef = nullptr;
    c.partial_cap = c.small_cap = c.coef_cap = 0;
    cudaStreamDestroy(c.stream);
    cudaStreamDestroy(c.comm_stream);
    c.stream = c.comm_stream = nullptr;
    c.rank = 0; c.nranks = 1;
    c.ready = false;
    return RAILS_OK;
}

extern "C" int rails_cuda_is_initialized(void) { return ctx().ready ? 1 : 0; }

extern "C" int rails_cuda_sync(void)
{
    RAILS_REQUIRE_READY();
    RAILS_CUDA_TRY(cudaStreamSynchronize(ctx().stream));
    return rails::check_exchange_error("rails_cuda_sync");
}

extern "C" const char *rails_cuda_last_error(void) { return ctx().last_error.c_str(); }
extern "C" long long rails_cuda_error_count(void) { return ctx().errors.load(); }
extern "C" long long rails_cuda_kernel_launches(void) { return ctx().launches.load(); }
extern "C" long long rails_cuda_host_syncs(void) { return ctx().host_syncs.load(); }
extern "C" long long rails_cuda_bytes_in_use(void) { return ctx().bytes_in_use.load(); }

extern "C" int rails_cuda_event_record(int slot)
{
    RAILS_REQUIRE_READY();
    rails::Context &c = ctx();
    if (slot < 0 || slot >= (int)c.events.size())
        return rails::fail(RAILS_ERR_ARG, "event slot %d out of range", slot);
    if (!c.events[slot])
        RAILS_CUDA_TRY(cudaEventCreate(&c.events[slot]));
    RAILS_CUDA_TRY(cudaEventRecord(c.events[slot], c.stream));
    return RAILS_OK;
}

extern "C" int rails_cuda_event_elapsed_ms(int a, int b, float *ms)
{
    RAILS_REQUIRE_READY();
    rails::Context &c = ctx();
    int n = (int)c.events.size();
    if (a < 0 || a >= n || b < 0 || b >= n || !c.events[a] || !c.events[b] || !ms)
        return rails::fail(RAILS_ERR_ARG, "bad event slots %d, %d", a, b);
    RAILS_CUDA_TRY(cudaEventSynchronize(c.events[b]));
    RAILS_CUDA_TRY(cudaEventElapsedTime(ms, c.events[a], c.events[b]));
    return RAILS_OK;
}

// ---- communicator -------------------------------------------------------

extern "C" int rails_cuda_comm_unique_id(char id[RAILS_NCCL_ID_BYTES])
{
    static_assert(sizeof(ncclUniqueId) <= RAILS_NCCL_ID_BYTES, "id size");
    ncclUniqueId uid;
    RAILS_NCCL_TRY(ncclGetUniqueId(&uid));
    memset(id, 0, RAILS_NCCL_ID_BYTES);
    memcpy(id, &uid, sizeof(uid));
    return RAILS_OK;
}

extern "C" int rails_cuda_comm_init(int rank, int nranks, const char id[RAILS_NCCL_ID_BYTES])
{
    RAILS_REQUIRE_READY();
    rails::Context &c = ctx();
    if (nranks < 1 || rank < 0 || rank >= nranks)
        return rails::fail(RAILS_ERR_ARG, "bad rank %d of %d", rank, nranks);
    rails::release_mailboxes();
    if (c.comm) { ncclCommDestroy(c.comm); c.comm = nullptr; }
    c.rank = rank;
    c.nranks = nranks;
    if (nranks == 1)
        return RAILS_OK;
    ncclUniqueId uid;
    memcpy(&uid, id, sizeof(uid));
    RAILS_NCCL_TRY(ncclCommInitRank(&c.comm, nranks, uid, rank));
    // NCCL connects lazily: pay for that here, not inside the first solve
    {
        double *w = nullptr;
        RAILS_CUDA_TRY(cudaMalloc(&w, 65536 * sizeof(double)));
        RAILS_CUDA_TRY(cudaMemset(w, 0, 65536 * sizeof(double)));
        for (size_t n : {(size_t)1, (size_t)256, (size_t)65536})
            RAILS_NCCL_TRY(ncclAllReduce(w, w, n, ncclDouble, ncclSum, c.comm, c.stream));
        RAILS_NCCL_TRY(ncclAllReduce(w, w, 1, ncclDouble, ncclMax, c.comm, c.stream));
        RAILS_CUDA_TRY(cudaStreamSynchronize(c.stream));
        cudaFree(w);
    }
    return rails::setup_mailboxes();
}

extern "C" int rails_cuda_comm_p2p(void) { return ctx().p2p ? 1 : 0; }
extern "C" int rails_cuda_comm_rank(void) { return ctx().rank; }
extern "C" int rails_cuda_comm_size(void) { return ctx().nranks; }

extern "C" int rails_partition_rows(long long m, int nranks, int rank,
                                    long long *begin, long long *end)
{
    if (nranks < 1 || rank < 0 || rank >= nranks || m < 0 || !begin || !end)
        return rails::fail(RAILS_ERR_ARG, "bad partition request");
    long long base = m / nranks, rem = m % nranks;
    *begin = rank * base + (rank < rem ? rank : rem);
    *end = *begin + base + (rank < rem ? 1 : 0);
    return RAILS_OK;
}

// ---- multivector storage ---------------------------------------------------

static int check_range(rails_mv_t mv, int col0, int ncols, const char *what)
{
    if (!mv)
        return rails::fail(RAILS_ERR_ARG, "%s: null multivector", what);
    if (col0 < 0 || ncols < 0 || col0 + ncols > mv->cols)
        return rails::fail(RAILS_ERR_ARG, "%s: columns [%d,%d) outside capacity %d",
                           what, col0, col0 + ncols, mv->cols);
    return RAILS_OK;
}

extern "C" int rails_mv_alloc(long long rows, int cols, int flags, rails_mv_t *out)
{
    RAILS_REQUIRE_READY();
    if (rows < 0 || cols < 0 || !out)
        return rails::fail(RAILS_ERR_ARG, "rails_mv_alloc(%lld, %d)", rows, cols);
    rails_mv_s *mv = new rails_mv_s();
    mv->rows = rows;
    mv->ld = ((rows + 15) / 16) * 16;
    if (mv->ld == 0) mv->ld = 16;
    mv->cols = cols;
    mv->flags = flags;
    mv->bytes = (size_t)mv->ld * (size_t)(cols > 0 ? cols : 1) * sizeof(double);
    void *block = nullptr;
    int rc = rails::dev_alloc(&block, mv->bytes);
    if (rc)
    {
        delete mv;
        return rc;
    }
    mv->data = static_cast<double *>(block);
    ctx().bytes_in_use.fetch_add((long long)mv->bytes);
    *out = mv;
    return RAILS_OK;
}

extern "C" int rails_mv_free(rails_mv_t mv)
{
    if (!mv)
        return RAILS_OK;
    if (ctx().ready && mv->data)
    {
        rails::dev_free(mv->data);
        ctx().bytes_in_use.fetch_sub((long long)mv->bytes);
    }
    delete mv;
    return RAILS_OK;
}

extern "C" long long rails_mv_rows(rails_mv_t mv) { return mv ? mv->rows : -1; }
extern "C" long long rails_mv_ld(rails_mv_t mv) { return mv ? mv->ld : -1; }
extern "C" int rails_mv_cols(rails_mv_t mv) { return mv ? mv->cols : -1; }
extern "C" int rails_mv_flags(rails_mv_t mv) { return mv ? mv->flags : -1; }

extern "C" int rails_mv_copy_cols(rails_mv_t dst, int dcol, rails_mv_t src, int scol, int ncols)
{
    RAILS_REQUIRE_READY();
    int rc = check_range(dst, dcol, ncols, "copy_cols dst");
    if (rc) return rc;
    rc = check_range(src, scol, ncols, "copy_cols src");
    if (rc) return rc;
    if (dst->rows != src->rows)
        return rails::fail(RAILS_ERR_ARG, "copy_cols: %lld vs %lld rows", dst->rows, src->rows);
    if (ncols == 0 || dst->rows == 0)
        return RAILS_OK;
    RAILS_CUDA_TRY(cudaMemcpy2DAsync(dst->data + (size_t)dcol * dst->ld, dst->ld * sizeof(double),
                                     src->data + (size_t)scol * src->ld, src->ld * sizeof(double),
                                     (size_t)dst->rows * sizeof(double), ncols,
                                     cudaMemcpyDeviceToDevice, ctx().stream));
    return RAILS_OK;
}

extern "C" int rails_mv_upload(rails_mv_t mv, int col0, int ncols, const double *host, long long ldh)
{
    RAILS_REQUIRE_READY();
    int rc = check_range(mv, col0, ncols, "upload");
    if (rc) return rc;
    if (ncols == 0 || mv->rows == 0)
        return RAILS_OK;
    if (!host || ldh < mv->rows)
        return rails::fail(RAILS_ERR_ARG, "upload: bad host block");
    if ((size_t)mv->rows * sizeof(double) >= (8u << 20))
    {
        // large columns: staged, pipelined copies column by column
        for (int j = 0; j < ncols; ++j)
        {
            rc = rails::copy_h2d(mv->data + (size_t)(col0 + j) * mv->ld, host + (size_t)j * ldh,
                                 (size_t)mv->rows * sizeof(double));
            if (rc) return rc;
        }
        return RAILS_OK;
    }
    RAILS_CUDA_TRY(cudaMemcpy2DAsync(mv->data + (size_t)col0 * mv->ld, mv->ld * sizeof(double),
                                     host, ldh * sizeof(double),
                                     (size_t)mv->rows * sizeof(double), ncols,
                                     cudaMemcpyHostToDevice, ctx().stream));
    // pageable host memory may be reused by the caller right after return
    RAILS_CUDA_TRY(cudaStreamSynchronize(ctx().stream));
    return RAILS_OK;
}

extern "C" int rails_mv_download(rails_mv_t mv, int col0, int ncols, double *host, long long ldh)
{
    RAILS_REQUIRE_READY();
    int rc = check_range(mv, col0, ncols, "download");
    if (rc) return rc;
    if (ncols == 0 || mv->rows == 0)
        return RAILS_OK;
    if (!host || ldh < mv->rows)
        return rails::fail(RAILS_ERR_ARG, "download: bad host block");
    return rails::copy_d2h_2d(host, (size_t)ldh * sizeof(double), mv->data + (size_t)col0 * mv->ld,
                              (size_t)mv->ld * sizeof(double), (size_t)mv->rows * sizeof(double), ncols);
}

extern "C" int rails_mv_get(rails_mv_t mv, long long row, int col, double *value)
{
    RAILS_REQUIRE_READY();
    int rc = check_range(mv, col, 1, "get");
    if (rc) return rc;
    if (row < 0 || row >= mv->rows || !value)
        return rails::fail(RAILS_ERR_ARG, "get: row %lld outside %lld", row, mv->rows);
    RAILS_CUDA_TRY(cudaMemcpyAsync(value, mv->data + (size_t)col * mv->ld + row, sizeof(double),
                                   cudaMemcpyDeviceToHost, ctx().stream));
    RAILS_CUDA_TRY(cudaStreamSynchronize(ctx().stream));
    return RAILS_OK;
}

extern "C" int rails_mv_set(rails_mv_t mv, long long row, int col, double value)
{
    RAILS_REQUIRE_READY();
    int rc = check_range(mv, col, 1, "set");
    if (rc) return rc;
    if (row < 0 || row >= mv->rows)
        return rails::fail(RAILS_ERR_ARG, "set: row %lld outside %lld", row, mv->rows);
    RAILS_CUDA_TRY(cudaMemcpyAsync(mv->data + (size_t)col * mv->ld + row, &value, sizeof(double),
                                   cudaMemcpyHostToDevice, ctx().stream));
    RAILS_CUDA_TRY(cudaStreamSynchronize(ctx().stream));
    return RAILS_OK;
}
