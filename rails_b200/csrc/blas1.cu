// BLAS-1 family, norm_inf and the bit-exact random() fill.
//
// All of these are pure HBM streams (fill/random 8 B per
// element, scal 16 B, axpy 24 B).  Columns start on 128-byte lines (ld is a
// multiple of 16 doubles), so every kernel moves 128-bit words; the padding
// rows below `rows` are never mixed into another row, which makes it safe for
// the elementwise kernels to run over the even-padded row count.
#include "internal.h"

using rails::ctx;

namespace
{

constexpr int EW_THREADS = 256;

__global__ void __launch_bounds__(EW_THREADS)
fill_kernel(double *__restrict__ x, long long ld, long long rows2, double value)
{
    double2 *col = reinterpret_cast<double2 *>(x + (size_t)blockIdx.y * ld);
    const double2 v = make_double2(value, value);
    for (long long i = (long long)blockIdx.x * EW_THREADS + threadIdx.x; i < rows2;
         i += (long long)gridDim.x * EW_THREADS)
        col[i] = v;
}

__global__ void __launch_bounds__(EW_THREADS)
scal_kernel(double *__restrict__ x, long long ld, long long rows2, double alpha)
{
    double2 *col = reinterpret_cast<double2 *>(x + (size_t)blockIdx.y * ld);
    for (long long i = (long long)blockIdx.x * EW_THREADS + threadIdx.x; i < rows2;
         i += (long long)gridDim.x * EW_THREADS)
    {
        double2 v = col[i];
        v.x *= alpha;
        v.y *= alpha;
        col[i] = v;
    }
}

// x *= 1 / sqrt(|*d|): `v /= v.norm()` of StlWrapper::orthogonalize (:311,319)
// with the squared norm still on the device.  1.0/norm and the product are
// separately rounded, as on the host.
__global__ void __launch_bounds__(EW_THREADS)
scal_inv_norm_kernel(double *__restrict__ x, long long ld, long long rows2,
                     const double *__restrict__ d)
{
    const double s = 1.0 / sqrt(fabs(*d));
    double2 *col = reinterpret_cast<double2 *>(x + (size_t)blockIdx.y * ld);
    for (long long i = (long long)blockIdx.x * EW_THREADS + threadIdx.x; i < rows2;
         i += (long long)gridDim.x * EW_THREADS)
    {
        double2 v = col[i];
        v.x *= s;
        v.y *= s;
        col[i] = v;
    }
}

__global__ void __launch_bounds__(EW_THREADS)
axpy_kernel(double *__restrict__ y, long long ldy, const double *__restrict__ x,
            long long ldx, long long rows2, double alpha)
{
    double2 *yc = reinterpret_cast<double2 *>(y + (size_t)blockIdx.y * ldy);
    const double2 *xc = reinterpret_cast<const double2 *>(x + (size_t)blockIdx.y * ldx);
    for (long long i = (long long)blockIdx.x * EW_THREADS + threadIdx.x; i < rows2;
         i += (long long)gridDim.x * EW_THREADS)
    {
        double2 a = xc[i];
        double2 b = yc[i];
        b.x = fma(alpha, a.x, b.x);
        b.y = fma(alpha, a.y, b.y);
        yc[i] = b;
    }
}

__global__ void __launch_bounds__(EW_THREADS)
norm_inf_kernel(const double *__restrict__ x, long long ld, long long rows, int ncols,
                unsigned long long *__restrict__ out)
{
    double best = 0.0;
    for (long long i = (long long)blockIdx.x * EW_THREADS + threadIdx.x; i < rows;
         i += (long long)gridDim.x * EW_THREADS)
    {
        double s = 0.0;
        for (int j = 0; j < ncols; ++j)
            s += fabs(x[i + (size_t)j * ld]);
        best = fmax(best, s);
    }
    for (int off = 16; off > 0; off >>= 1)
        best = fmax(best, __shfl_xor_sync(0xffffffffu, best, off));
    // non-negative doubles order like their bit patterns
    if ((threadIdx.x & 31) == 0)
        atomicMax(out, (unsigned long long)__double_as_longlong(best));
}

// ---- minstd_rand0 with jump-ahead ----------------------------------------
// libstdc++: default_random_engine == minstd_rand0, x <- 16807 x mod (2^31-1);
// generate_canonical<double,53> takes two draws x1, x2:
//     u = ((x1-1) + (x2-1) * R) / (R*R),  R = 2147483646   (all in double,
//     each operation rounded), and uniform_real_distribution(-1,1) returns
//     u * 2 + (-1).  Element (i, j) of an m x n fill (row-major order,
//     src/StlWrapper.cpp:419-421) consumes draws 2(i n + j) and 2(i n + j)+1.

constexpr unsigned long long LCG_A = 16807ull;
constexpr unsigned long long LCG_P = 2147483647ull;

__device__ __forceinline__ unsigned long long mulmod31(unsigned long long a, unsigned long long b)
{
    unsigned long long v = a * b;               // < 2^62
    v = (v & LCG_P) + (v >> 31);
    v = (v & LCG_P) + (v >> 31);
    return v >= LCG_P ? v - LCG_P : v;
}

__device__ __forceinline__ unsigned long long powmod31(unsigned long long base, unsigned long long e)
{
    unsigned long long r = 1;
    while (e)
    {
        if (e & 1ull)
            r = mulmod31(r, base);
        base = mulmod31(base, base);
        e >>= 1;
    }
    return r;
}

// Thread t fills rows t, t + stride, ...: one jump-ahead power per thread for
// its first row, then the state at the start of the next row is the current
// row start times `jump` = a^(2 * ncols * stride) (computed on the host).
__global__ void __launch_bounds__(EW_THREADS)
random_kernel(double *__restrict__ x, long long ld, long long rows, int ncols,
              long long row_offset, unsigned long long x0, unsigned long long jump)
{
    const double R = 2147483646.0;
    const double R2 = 4611686009837453316.0;   // double(R*R), exact product rounded
    long long i = (long long)blockIdx.x * EW_THREADS + threadIdx.x;
    if (i >= rows)
        return;
    const unsigned long long first = 2ull * (unsigned long long)(row_offset + i) * (unsigned long long)ncols;
    // state after `first` draws
    unsigned long long row_state = mulmod31(powmod31(LCG_A, first), x0);
    for (; i < rows; i += (long long)gridDim.x * EW_THREADS)
    {
        unsigned long long s = row_state;
        for (int j = 0; j < ncols; ++j)
        {
            s = mulmod31(s, LCG_A);
            double d1 = (double)(s - 1ull);
            s = mulmod31(s, LCG_A);
            double d2 = (double)(s - 1ull);
            double sum = __dadd_rn(d1, __dmul_rn(d2, R));
            double u = __ddiv_rn(sum, R2);
            if (u >= 1.0)
                u = 0.99999999999999989;        // nextafter(1, 0), as libstdc++
            x[i + (size_t)j * ld] = __dadd_rn(__dmul_rn(u, 2.0), -1.0);
        }
        row_state = mulmod31(row_state, jump);
    }
}

unsigned long long host_mulmod31(unsigned long long a, unsigned long long b)
{
    return (unsigned long long)((unsigned __int128)a * b % LCG_P);
}

unsigned long long host_powmod31(unsigned long long base, unsigned long long e)
{
    unsigned long long r = 1;
    base %= LCG_P;
    while (e)
    {
        if (e & 1ull) r = host_mulmod31(r, base);
        base = host_mulmod31(base, base);
        e >>= 1;
    }
    return r;
}

int ew_grid(long long work)
{
    long long b = (work + EW_THREADS - 1) / EW_THREADS;
    long long cap = (long long)ctx().sm_count * 8;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return (int)b;
}

int check(rails_mv_t mv, int col0, int ncols, const char *what)
{
    if (!mv)
        return rails::fail(RAILS_ERR_ARG, "%s: null multivector", what);
    if (col0 < 0 || ncols < 0 || col0 + ncols > mv->cols)
        return rails::fail(RAILS_ERR_ARG, "%s: columns [%d,%d) outside capacity %d", what,
                           col0, col0 + ncols, mv->cols);
    return RAILS_OK;
}

}

namespace rails
{
// used by the Gram-Schmidt chain in dot.cu
int scal_inv_norm_device(double *x, long long ld, long long rows, const double *d)
{
    long long rows2 = (rows + 1) / 2;
    if (rows2 == 0) return RAILS_OK;
    dim3 grid(ew_grid(rows2), 1);
    RAILS_LAUNCH(scal_inv_norm_kernel, grid, EW_THREADS, 0, x, ld, rows2, d);
    RAILS_CHECK_LAUNCH();
    return RAILS_OK;
}
}

extern "C" int rails_mv_fill(rails_mv_t mv, int col0, int ncols, double value)
{
    RAILS_REQUIRE_READY();
    int rc = check(mv, col0, ncols, "fill");
    if (rc) return rc;
    long long rows2 = (mv->rows + 1) / 2;
    if (ncols == 0 || rows2 == 0) return RAILS_OK;
    dim3 grid(ew_grid(rows2), ncols);
    RAILS_LAUNCH(fill_kernel, grid, EW_THREADS, 0, mv->data + (size_t)col0 * mv->ld, mv->ld, rows2, value);
    RAILS_CHECK_LAUNCH();
    return RAILS_OK;
}

extern "C" int rails_mv_scal(rails_mv_t mv, int col0, int ncols, double alpha)
{
    RAILS_REQUIRE_READY();
    int rc = check(mv, col0, ncols, "scal");
    if (rc) return rc;
    long long rows2 = (mv->rows + 1) / 2;
    if (ncols == 0 || rows2 == 0) return RAILS_OK;
    dim3 grid(ew_grid(rows2), ncols);
    RAILS_LAUNCH(scal_kernel, grid, EW_THREADS, 0, mv->data + (size_t)col0 * mv->ld, mv->ld, rows2, alpha);
    RAILS_CHECK_LAUNCH();
    return RAILS_OK;
}

extern "C" int rails_mv_axpy(rails_mv_t y, int yc0, double alpha, rails_mv_t x, int xc0, int ncols)
{
    RAILS_REQUIRE_READY();
    int rc = check(y, yc0, ncols, "axpy y");
    if (rc) return rc;
    rc = check(x, xc0, ncols, "axpy x");
    if (rc) return rc;
    if (x->rows != y->rows)
        return rails::fail(RAILS_ERR_ARG, "axpy: %lld vs %lld rows", x->rows, y->rows);
    long long rows2 = (y->rows + 1) / 2;
    if (ncols == 0 || rows2 == 0) return RAILS_OK;
    dim3 grid(ew_grid(rows2), ncols);
    RAILS_LAUNCH(axpy_kernel, grid, EW_THREADS, 0, y->data + (size_t)yc0 * y->ld, y->ld,
                 x->data + (size_t)xc0 * x->ld, x->ld, rows2, alpha);
    RAILS_CHECK_LAUNCH();
    return RAILS_OK;
}

extern "C" int rails_mv_norm_inf(rails_mv_t mv, int col0, int ncols, double *out)
{
    RAILS_REQUIRE_READY();
    int rc = check(mv, col0, ncols, "norm_inf");
    if (rc) return rc;
    if (!out) return rails::fail(RAILS_ERR_ARG, "norm_inf: null result");
    rails::Context &c = ctx();
    *out = 0.0;
    if (ncols == 0 || mv->rows == 0) return RAILS_OK;
    unsigned long long *slot = reinterpret_cast<unsigned long long *>(c.d_small);
    RAILS_CUDA_TRY(cudaMemsetAsync(slot, 0, sizeof(unsigned long long), c.stream));
    RAILS_LAUNCH(norm_inf_kernel, ew_grid(mv->rows), EW_THREADS, 0,
                 mv->data + (size_t)col0 * mv->ld, mv->ld, mv->rows, ncols, slot);
    RAILS_CHECK_LAUNCH();
    if (c.comm && mv->flags == RAILS_MV_DISTRIBUTED)
        RAILS_NCCL_TRY(ncclAllReduce(c.d_small, c.d_small, 1, ncclDouble, ncclMax, c.comm, c.stream));
    RAILS_CUDA_TRY(cudaMemcpyAsync(c.h_small, c.d_small, sizeof(double), cudaMemcpyDeviceToHost, c.stream));
    RAILS_CUDA_TRY(cudaStreamSynchronize(c.stream));
    { int xrc = rails::check_exchange_error("blas1"); if (xrc) return xrc; }
    *out = c.h_small[0];
    return RAILS_OK;
}

extern "C" int rails_mv_random(rails_mv_t mv, int col0, int ncols, long long row_offset,
                               unsigned int seed)
{
    RAILS_REQUIRE_READY();
    int rc = check(mv, col0, ncols, "random");
    if (rc) return rc;
    if (ncols == 0 || mv->rows == 0) return RAILS_OK;
    // linear_congruential_engine(seed): state = seed mod p, 0 -> 1
    unsigned long long x0 = (unsigned long long)seed % LCG_P;
    if (x0 == 0) x0 = 1;
    const int grid = ew_grid(mv->rows);
    const unsigned long long stride = (unsigned long long)grid * EW_THREADS;
    const unsigned long long jump = host_powmod31(LCG_A, 2ull * (unsigned long long)ncols * stride);
    RAILS_LAUNCH(random_kernel, grid, EW_THREADS, 0,
                 mv->data + (size_t)col0 * mv->ld, mv->ld, mv->rows, ncols, row_offset, x0, jump);
    RAILS_CHECK_LAUNCH();
    return RAILS_OK;
}

namespace
{

__global__ void __launch_bounds__(EW_THREADS)
scaled_copy_kernel(double *__restrict__ y, long long ldy, const double *__restrict__ x,
                   long long ldx, long long rows2, double alpha)
{
    double2 *yc = reinterpret_cast<double2 *>(y + (size_t)blockIdx.y * ldy);
    const double2 *xc = reinterpret_cast<const double2 *>(x + (size_t)blockIdx.y * ldx);
    for (long long i = (long long)blockIdx.x * EW_THREADS + threadIdx.x; i < rows2;
         i += (long long)gridDim.x * EW_THREADS)
    {
        double2 a = xc[i];
        a.x *= alpha;
        a.y *= alpha;
        yc[i] = a;
    }
}

__global__ void __launch_bounds__(EW_THREADS)
add_scaled_kernel(double *__restrict__ y, long long ldy, const double *__restrict__ x,
                  long long ldx, long long rows2, double sign, double alpha)
{
    double2 *yc = reinterpret_cast<double2 *>(y + (size_t)blockIdx.y * ldy);
    const double2 *xc = reinterpret_cast<const double2 *>(x + (size_t)blockIdx.y * ldx);
    for (long long i = (long long)blockIdx.x * EW_THREADS + threadIdx.x; i < rows2;
         i += (long long)gridDim.x * EW_THREADS)
    {
        double2 a = xc[i];
        double2 b = yc[i];
        // round(alpha*x) first, then add: same bits as the two-step sequence
        b.x = __dadd_rn(b.x, sign * __dmul_rn(alpha, a.x));
        b.y = __dadd_rn(b.y, sign * __dmul_rn(alpha, a.y));
        yc[i] = b;
    }
}

}

extern "C" int rails_mv_scaled_copy(rails_mv_t dst, int dcol, double alpha, rails_mv_t src, int scol, int ncols)
{
    RAILS_REQUIRE_READY();
    int rc = check(dst, dcol, ncols, "scaled_copy dst");
    if (rc) return rc;
    rc = check(src, scol, ncols, "scaled_copy src");
    if (rc) return rc;
    if (src->rows != dst->rows)
        return rails::fail(RAILS_ERR_ARG, "scaled_copy: %lld vs %lld rows", src->rows, dst->rows);
    long long rows2 = (dst->rows + 1) / 2;
    if (ncols == 0 || rows2 == 0) return RAILS_OK;
    dim3 grid(ew_grid(rows2), ncols);
    RAILS_LAUNCH(scaled_copy_kernel, grid, EW_THREADS, 0, dst->data + (size_t)dcol * dst->ld, dst->ld,
                 src->data + (size_t)scol * src->ld, src->ld, rows2, alpha);
    RAILS_CHECK_LAUNCH();
    return RAILS_OK;
}

extern "C" int rails_mv_add_scaled(rails_mv_t y, int yc0, double sign, double alpha, rails_mv_t x,
                                   int xc0, int ncols)
{
    RAILS_REQUIRE_READY();
    int rc = check(y, yc0, ncols, "add_scaled y");
    if (rc) return rc;
    rc = check(x, xc0, ncols, "add_scaled x");
    if (rc) return rc;
    if (x->rows != y->rows)
        return rails::fail(RAILS_ERR_ARG, "add_scaled: %lld vs %lld rows", x->rows, y->rows);
    if (sign != 1.0 && sign != -1.0)
        return rails::fail(RAILS_ERR_ARG, "add_scaled: sign must be +1 or -1");
    long long rows2 = (y->rows + 1) / 2;
    if (ncols == 0 || rows2 == 0) return RAILS_OK;
    dim3 grid(ew_grid(rows2), ncols);
    RAILS_LAUNCH(add_scaled_kernel, grid, EW_THREADS, 0, y->data + (size_t)yc0 * y->ld, y->ld,
                 x->data + (size_t)xc0 * x->ld, x->ld, rows2, sign, alpha);
    RAILS_CHECK_LAUNCH();
    return RAILS_OK;
}
