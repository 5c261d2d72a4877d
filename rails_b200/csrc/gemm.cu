// Tall-skinny times small:  Out = alpha * V * C + beta * Out
// (V is m x k in HBM, C is k x c, c small).
//
// Replaces StlWrapper::operator* (src/StlWrapper.cpp:168-187, dgemm 'N','N')
// and EpetraMV::operator* (src/Epetra_MultiVectorWrapper.cpp:194-241) for the
// shapes the solver produces: c = 1 in resid_lanczos and CGS2
// (src/LyapunovSolver.hpp:396,402), c = L for the Ritz vectors (:443), c = r
// at a restart (:265,290).  For c = 1 this is a GEMV with 0.25 flop/B: the
// roofline is the HBM stream of 8 m (k + c) bytes (+ 8 m c when beta != 0).
//
// Each thread owns 2*RP rows and CT output columns; V is streamed column by
// column with 128-bit loads (8 columns in flight), the coefficient block sits
// in shared memory and is broadcast.  The accumulation over the k columns is
// sequential with fma, and alpha/beta are applied after the sum, so
// `y += V*z` gives the same bits whether or not the product is materialised.
#include "internal.h"

using rails::ctx;

namespace
{

constexpr int GEMM_THREADS = 256;
constexpr int GEMM_KCHUNK = 256;      // coefficient rows staged per pass
constexpr int GEMM_UNROLL = 8;

template<int RP, int CT>
__global__ void __launch_bounds__(GEMM_THREADS)
gemm_kernel(const double *__restrict__ V, long long ldv, int k,
            const double *__restrict__ C, int ldc,
            double alpha, double beta, double *__restrict__ Out, long long ldo,
            long long rows2)
{
    __shared__ double cs[GEMM_KCHUNK * CT];
    // row pairs handled by this thread
    long long r2[RP];
#pragma unroll
    for (int v = 0; v < RP; ++v)
        r2[v] = ((long long)blockIdx.x * RP + v) * GEMM_THREADS + threadIdx.x;

    double2 acc[CT][RP];
#pragma unroll
    for (int t = 0; t < CT; ++t)
#pragma unroll
        for (int v = 0; v < RP; ++v)
            acc[t][v] = make_double2(0.0, 0.0);

    const double *Cb = C + (size_t)blockIdx.y * CT * ldc;
    for (int j0 = 0; j0 < k; j0 += GEMM_KCHUNK)
    {
        const int kc = min(GEMM_KCHUNK, k - j0);
        __syncthreads();
        for (int i = threadIdx.x; i < kc * CT; i += GEMM_THREADS)
        {
            const int j = i / CT, t = i - j * CT;
            cs[i] = C ? Cb[(size_t)t * ldc + j0 + j] : 0.0;
        }
        __syncthreads();
        const double2 *Vc = reinterpret_cast<const double2 *>(V + (size_t)j0 * ldv);
        const long long ldv2 = ldv / 2;
        int j = 0;
        for (; j + GEMM_UNROLL <= kc; j += GEMM_UNROLL)
        {
            double2 x[GEMM_UNROLL][RP];
#pragma unroll
            for (int u = 0; u < GEMM_UNROLL; ++u)
#pragma unroll
                for (int v = 0; v < RP; ++v)
                    x[u][v] = (r2[v] < rows2) ? __ldg(Vc + (size_t)(j + u) * ldv2 + r2[v])
                                              : make_double2(0.0, 0.0);
#pragma unroll
            for (int u = 0; u < GEMM_UNROLL; ++u)
#pragma unroll
                for (int t = 0; t < CT; ++t)
                {
                    const double cf = cs[(j + u) * CT + t];
#pragma unroll
                    for (int v = 0; v < RP; ++v)
                    {
                        acc[t][v].x = fma(x[u][v].x, cf, acc[t][v].x);
                        acc[t][v].y = fma(x[u][v].y, cf, acc[t][v].y);
                    }
                }
        }
        for (; j < kc; ++j)
        {
#pragma unroll
            for (int v = 0; v < RP; ++v)
            {
                const double2 x = (r2[v] < rows2) ? __ldg(Vc + (size_t)j * ldv2 + r2[v])
                                                  : make_double2(0.0, 0.0);
#pragma unroll
                for (int t = 0; t < CT; ++t)
                {
                    const double cf = cs[j * CT + t];
                    acc[t][v].x = fma(x.x, cf, acc[t][v].x);
                    acc[t][v].y = fma(x.y, cf, acc[t][v].y);
                }
            }
        }
    }

#pragma unroll
    for (int t = 0; t < CT; ++t)
    {
        double2 *oc = reinterpret_cast<double2 *>(Out + ((size_t)blockIdx.y * CT + t) * ldo);
#pragma unroll
        for (int v = 0; v < RP; ++v)
        {
            if (r2[v] >= rows2) continue;
            double2 o;
            if (beta == 0.0)
            {
                o.x = alpha * acc[t][v].x;
                o.y = alpha * acc[t][v].y;
            }
            else
            {
                const double2 old = oc[r2[v]];
                o.x = fma(beta, old.x, alpha * acc[t][v].x);
                o.y = fma(beta, old.y, alpha * acc[t][v].y);
            }
            oc[r2[v]] = o;
        }
    }
}

template<int RP, int CT>
int launch_gemm(const double *V, long long ldv, int k, const double *dC, int ldc, int ngroups,
                double alpha, double beta, double *Out, long long ldo, long long rows)
{
    const long long rows2 = (rows + 1) / 2;
    const long long per_block = (long long)GEMM_THREADS * RP;
    dim3 grid((unsigned)((rows2 + per_block - 1) / per_block), ngroups);
    auto kern = gemm_kernel<RP, CT>;
    RAILS_LAUNCH(kern, grid, GEMM_THREADS, 0, V, ldv, k, dC, ldc, alpha, beta, Out, ldo, rows2);
    RAILS_CHECK_LAUNCH();
    return RAILS_OK;
}

}

namespace rails
{

// dC: device pointer to the k x c coefficient block (column-major, ldc).
int gemm_device(const double *V, long long ldv, int k, const double *dC, int ldc, int c,
                double alpha, double beta, double *Out, long long ldo, long long rows)
{
    if (c == 0 || rows == 0)
        return RAILS_OK;
    if (k == 0)
        dC = nullptr;          // empty sum: Out = beta * Out
    // many result columns: FP64-arithmetic-bound, V read once (gram.cu)
    if (dC && gemm_dmma_fits(k, c) && (size_t)V % 16 == 0 && (size_t)Out % 16 == 0 && ldv % 2 == 0 &&
        ldo % 2 == 0)
        return gemm_dmma_device(V, ldv, k, dC, ldc, c, alpha, beta, Out, ldo, rows);
    int t = 0;
    while (t < c)
    {
        const int left = c - t;
        const double *Ct = dC ? dC + (size_t)t * ldc : nullptr;
        double *Ot = Out + (size_t)t * ldo;
        int rc;
        if (left >= 8)      { int g = left / 8; rc = launch_gemm<1, 8>(V, ldv, k, Ct, ldc, g, alpha, beta, Ot, ldo, rows); t += 8 * g; }
        else if (left >= 4) { rc = launch_gemm<1, 4>(V, ldv, k, Ct, ldc, 1, alpha, beta, Ot, ldo, rows); t += 4; }
        else if (left >= 2) { rc = launch_gemm<2, 2>(V, ldv, k, Ct, ldc, 1, alpha, beta, Ot, ldo, rows); t += 2; }
        else                { rc = launch_gemm<2, 1>(V, ldv, k, Ct, ldc, 1, alpha, beta, Ot, ldo, rows); t += 1; }
        if (rc) return rc;
    }
    return RAILS_OK;
}

}

namespace
{

int check_gemm(rails_mv_t V, int vc0, int k, int c, rails_mv_t Out, int oc0)
{
    if (!V || !Out)
        return rails::fail(RAILS_ERR_ARG, "gemm: null multivector");
    if (vc0 < 0 || k < 0 || vc0 + k > V->cols || oc0 < 0 || c < 0 || oc0 + c > Out->cols)
        return rails::fail(RAILS_ERR_ARG, "gemm: column range outside capacity");
    if (V->rows != Out->rows)
        return rails::fail(RAILS_ERR_ARG, "gemm: %lld vs %lld rows", V->rows, Out->rows);
    if (V->data == Out->data && k > 0 && c > 0 && oc0 < vc0 + k && vc0 < oc0 + c)
        return rails::fail(RAILS_ERR_ARG, "gemm: result overlaps the columns being read");
    return RAILS_OK;
}

}

extern "C" int rails_mv_gemm(rails_mv_t V, int vc0, int k, const double *C, int ldc, int c,
                             double alpha, double beta, rails_mv_t Out, int oc0)
{
    RAILS_REQUIRE_READY();
    int rc = check_gemm(V, vc0, k, c, Out, oc0);
    if (rc) return rc;
    if (c == 0) return RAILS_OK;
    if (k > 0 && (!C || ldc < k))
        return rails::fail(RAILS_ERR_ARG, "gemm: bad coefficient block");
    const double *dC = nullptr;
    if (k > 0)
    {
        rc = rails::upload_small(C, k, c, ldc, &dC);
        if (rc) return rc;
    }
    return rails::gemm_device(V->data + (size_t)vc0 * V->ld, V->ld, k, dC, k, c, alpha, beta,
                              Out->data + (size_t)oc0 * Out->ld, Out->ld, V->rows);
}

extern "C" int rails_mv_gemm_dev(rails_mv_t V, int vc0, int k, rails_mv_t C, int cc0, int c,
                                 double alpha, double beta, rails_mv_t Out, int oc0)
{
    RAILS_REQUIRE_READY();
    int rc = check_gemm(V, vc0, k, c, Out, oc0);
    if (rc) return rc;
    if (c == 0) return RAILS_OK;
    if (!C || cc0 < 0 || cc0 + c > C->cols || C->rows != k)
        return rails::fail(RAILS_ERR_ARG, "gemm_dev: coefficient block must be %d x %d", k, c);
    return rails::gemm_device(V->data + (size_t)vc0 * V->ld, V->ld, k,
                              C->data + (size_t)cc0 * C->ld, (int)C->ld, c, alpha, beta,
                              Out->data + (size_t)oc0 * Out->ld, Out->ld, V->rows);
}
