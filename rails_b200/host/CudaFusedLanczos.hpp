// Fused residual-Lanczos recurrence for the CUDA wrappers: the specialisation
// of RAILS::b200::FusedLanczos that the product loop picks up
// (RailsSolverImpl.hpp, resid_lanczos).  Same recurrence as
// /root/reference/src/LyapunovSolver.hpp:384-434, three kernels per step
// (rails_mv_multi_dot, rails_mv_multi_gemv, rails_mv_lanczos_update; see
// rails_b200/csrc/lanczos.cu) instead of ~19 wrapper calls:
//
//   g = [V AV B]' q                      (the `q /= beta` of the previous step
//                                         is applied on the fly and written back)
//   next = V (T AV'q) + AV (T V'q) + B (B'q);   alpha = next' q
//   next -= alpha q;  next -= beta q_prev;      beta = |next|
//
// The k x k products T * (V'q), T * (AV'q) stay on the host, as in the
// reference.  Set RAILS_FUSED_LANCZOS=0 to run the operation-by-operation
// sequence instead (A/B comparisons in the tests).
#ifndef RAILS_B200_CUDAFUSEDLANCZOS_HPP
#define RAILS_B200_CUDAFUSEDLANCZOS_HPP

#include "CudaMultiVectorWrapper.hpp"
#include "HostDenseMatrix.hpp"
#include "RailsSolver.hpp"

#include <cstdlib>
#include <cstring>
#include <iostream>
#include <vector>

namespace RAILS
{
namespace b200
{

template<>
struct BackendHealth<CudaMultiVectorWrapper>
{
    static long long errors() { return rails_cuda_error_count(); }
    static void fence()
    {
        static const bool profiling = std::getenv("RAILS_PROFILE") && std::atoi(std::getenv("RAILS_PROFILE")) != 0;
        if (profiling)
            rails_cuda_sync();
    }
};

// The projection dots with one host round trip (rails_mv_dot_batch).
template<>
struct BatchedDots<CudaMultiVectorWrapper, HostDenseMatrix>
{
    static void run(std::vector<std::pair<CudaMultiVectorWrapper const *, CudaMultiVectorWrapper const *> > const &pairs,
                    std::vector<HostDenseMatrix> &out)
    {
        const int n = (int)pairs.size();
        out.clear();
        std::vector<rails_mv_t> X(n), Y(n);
        std::vector<int> xc(n), a(n), yc(n), b(n), ld(n);
        std::vector<double *> ptr(n);
        bool ok = true;
        for (int i = 0; i < n; ++i)
        {
            CudaMultiVectorWrapper const &x = *pairs[i].first, &y = *pairs[i].second;
            out.push_back(HostDenseMatrix(x.N(), y.N()));
            X[i] = x.device_handle(); Y[i] = y.device_handle();
            xc[i] = x.device_col0(); yc[i] = y.device_col0();
            a[i] = x.N(); b[i] = y.N();
            if (x.M() != y.M() || ((!X[i] || !Y[i]) && a[i] > 0 && b[i] > 0))
                ok = false;
            if (!X[i] || !Y[i]) { a[i] = 0; b[i] = 0; X[i] = X[i] ? X[i] : Y[i]; Y[i] = Y[i] ? Y[i] : X[i]; }
        }
        for (int i = 0; i < n; ++i)
        {
            ptr[i] = out[i];
            ld[i] = std::max(out[i].LDA(), 1);
            if (!X[i]) ok = false;
        }
        if (!ok)
        {
            // irregular operands: the plain sequence reports the problem as dot() does
            out.clear();
            for (int i = 0; i < n; ++i)
                out.push_back(pairs[i].first->dot(*pairs[i].second));
            return;
        }
        int rc = rails_mv_dot_batch(n, X.data(), xc.data(), a.data(), Y.data(), yc.data(), b.data(),
                                    ptr.data(), ld.data());
        if (rc)
            std::cerr << "BatchedDots failed (" << rc << "): " << rails_cuda_last_error() << std::endl;
    }
};

template<>
struct FusedLanczos<CudaMultiVectorWrapper, HostDenseMatrix>
{
    static const bool enabled = true;

    // RAILS_FUSED_LANCZOS: unset / "1" = the device-resident chain
    // (rails_mv_lanczos_run: no host round trip inside the recurrence),
    // "step" = three fused kernels per step with the k x k products on the host,
    // "0" = the operation-by-operation sequence of the reference.
    static bool active()
    {
        const char *e = std::getenv("RAILS_FUSED_LANCZOS");
        return !(e && std::strcmp(e, "0") == 0);
    }
    static bool fits(int k, int n) { return 2 * k + n <= rails_mv_lanczos_max_columns(); }
    static bool per_step()
    {
        const char *e = std::getenv("RAILS_FUSED_LANCZOS");
        return e && std::strcmp(e, "step") == 0;
    }

    static void report(int rc, const char *what)
    {
        if (rc)
            std::cerr << "FusedLanczos::" << what << " failed (" << rc << "): "
                      << rails_cuda_last_error() << std::endl;
    }

    // First projection of the chain, enqueued before T exists (rails_mv_lanczos_begin)
    static void begin(CudaMultiVectorWrapper const &AV, CudaMultiVectorWrapper const &V,
                      CudaMultiVectorWrapper const &B, CudaMultiVectorWrapper &Q, int max_iter)
    {
        if (per_step() || AV.N() != V.N())
            return;
        Q.resize(max_iter + 1);
        rails_mv_t blocks[3] = {V.device_handle(), AV.device_handle(), B.device_handle()};
        int col0[3] = {V.device_col0(), AV.device_col0(), B.device_col0()};
        int ncols[3] = {V.N(), AV.N(), B.N()};
        report(rails_mv_lanczos_begin(3, blocks, col0, ncols, Q.device_handle(), Q.device_col0()),
               "lanczos_begin");
        Q.resize(1);
    }

    // Q holds the un-normalised random start vector in column 0 and has room
    // for max_iter + 1 columns.  Returns the number of Lanczos steps.
    static int run(CudaMultiVectorWrapper const &AV, CudaMultiVectorWrapper const &V,
                   HostDenseMatrix const &T, CudaMultiVectorWrapper const &B,
                   HostDenseMatrix &H, CudaMultiVectorWrapper &Q, int max_iter)
    {
        const int k = V.N(), n = B.N();
        if (AV.N() != k || T.M() != k || T.N() != k)
        {
            std::cerr << "FusedLanczos: inconsistent sizes" << std::endl;
            return 0;
        }
        Q.resize(max_iter + 1);
        rails_mv_t hq = Q.device_handle();
        const int q0 = Q.device_col0();
        rails_mv_t blocks[3] = {V.device_handle(), AV.device_handle(), B.device_handle()};
        int col0[3] = {V.device_col0(), AV.device_col0(), B.device_col0()};
        int ncols[3] = {k, k, n};

        if (!per_step())
        {
            std::vector<double> a(max_iter), b(max_iter);
            int done = 0;
            report(rails_mv_lanczos_run(3, blocks, col0, ncols, T, T.LDA(), hq, q0, max_iter, a.data(),
                                        b.data(), &done), "lanczos_run");
            for (int i = 0; i < done; ++i)
            {
                H(i, i) = a[i];
                if (b[i] < 1e-14)
                    break;                      // breakdown: no off-diagonal entry (:419-427)
                H(i + 1, i) = b[i];
                H(i, i + 1) = b[i];
                if (std::getenv("RAILS_DEBUG_LANCZOS"))
                    std::cerr << "lanczos chain step " << i << " alpha " << a[i] << " beta " << b[i] << std::endl;
            }
            return done;
        }

        // `Q.view(0) /= Q.norm()` (:375): the reciprocal is rounded first, the
        // scaling itself happens inside the first multi_dot
        Q.resize(1);
        double qscale = 1.0 / Q.norm();
        Q.resize(max_iter + 1);

        std::vector<double> g(2 * k + n), coef(2 * k + n);
        HostDenseMatrix Z(k, 1);
        double alpha = 0.0, beta = 0.0;
        int steps = 0;
        for (int i = 0; i < max_iter; i++)
        {
            const int cq = q0 + steps, cn = q0 + steps + 1;
            report(rails_mv_multi_dot(3, blocks, col0, ncols, hq, cq, qscale, g.data()), "multi_dot");

            // coefficient of V is T * (AV'q), of AV is T * (V'q), of B is B'q
            for (int j = 0; j < k; ++j) Z(j, 0) = g[j];
            HostDenseMatrix TZ = T * Z;
            for (int j = 0; j < k; ++j) coef[k + j] = TZ(j, 0);
            for (int j = 0; j < k; ++j) Z(j, 0) = g[k + j];
            TZ = T * Z;
            for (int j = 0; j < k; ++j) coef[j] = TZ(j, 0);
            for (int j = 0; j < n; ++j) coef[2 * k + j] = g[2 * k + j];

            report(rails_mv_multi_gemv(3, blocks, col0, ncols, coef.data(), hq, cn, hq, cq, &alpha),
                   "multi_gemv");
            H(steps, steps) = alpha;

            double norm = 0.0;
            report(rails_mv_lanczos_update(hq, cn, alpha, hq, cq, beta,
                                           steps > 0 ? hq : (rails_mv_t) nullptr, cq - 1, &norm),
                   "lanczos_update");
            beta = norm;
            if (std::getenv("RAILS_DEBUG_LANCZOS"))
                std::cerr << "lanczos fused step " << steps << " alpha " << alpha << " beta " << beta
                          << " g0 " << g[0] << " gB " << g[2 * k] << " qscale " << qscale << std::endl;
            if (beta < 1e-14)
            {
                steps++;
                break;
            }
            H(steps + 1, steps) = beta;
            H(steps, steps + 1) = beta;
            qscale = 1.0 / beta;         // applied by the next step's multi_dot
            steps++;
        }
        // the last `next` is dropped by the caller (Q.resize(steps)), so its
        // pending scale is never needed
        return steps;
    }
};

}
}

#endif
