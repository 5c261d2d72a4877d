// TEST INFRASTRUCTURE (oracle/_ref build): the UNMODIFIED reference solver
// loop (/root/reference/src/LyapunovSolver.hpp, with its own
// MatrixOrMultiVectorWrapper, StlTools, SlicotWrapper and LapackWrapper)
// instantiated with the PRODUCT's CUDA wrapper classes:
//
//     RAILS::Solver<CudaOperatorWrapper, CudaMultiVectorWrapper, HostDenseMatrix>
//
// This is the drop-in proof: the three classes of rails_b200/host/ satisfy the
// reference's plugin surface (src/LyapunovSolverDecl.hpp:9-51) without touching
// a line of the reference.  Compiled by oracle/build_ref.py from the sources
// where they lie under /root/reference; only the .so (oracle/_ref/) travels to
// the GPU box, where tests/test_gpu_dropin.py runs it next to the product loop.
#include <algorithm>
#include <chrono>
#include <iostream>
#include <utility>

#include "capture.hpp"

#include "CudaMultiVectorWrapper.hpp"
#include "CudaOperatorWrapper.hpp"
#include "HostDenseMatrix.hpp"

#include "src/LyapunovSolver.hpp"

using namespace RAILS;

extern "C" void rails_host_limit_blas_threads(void);

extern "C" int refloop_cuda_solve(
    int m, const int *rowptr, const int *colidx, const double *vals,
    int n_b, const double *B_in, unsigned seed,
    int nparams, const char *const *keys, const double *pvals,
    int hist_cap, double *hist_abs, double *hist_rel, int *hist_len,
    int *k_out, int v_cap_cols, double *V_out, double *T_out,
    double *seconds)
{
    if (!rails_cuda_is_initialized())
        return 200;
    rails_host_limit_blas_threads();      // same k x k BLAS threading as the product entry
    std::srand(seed);

    CudaOperatorWrapper A(m, m, rowptr, colidx, vals);
    CudaMultiVectorWrapper B(m, n_b);
    if (B_in)
        rails_mv_upload(B.device_handle(), 0, n_b, B_in, m);
    else
        B.random();

    // M is stored and never read by the reference loop
    Solver<CudaOperatorWrapper, CudaMultiVectorWrapper, HostDenseMatrix> solver(A, B, A);
    rails_oracle::ParameterList params(nparams, keys, pvals);
    int ret = solver.set_parameters(params);
    if (ret)
        return 100 + ret;

    CudaMultiVectorWrapper V(m, 1);
    HostDenseMatrix T;

    std::string log;
    {
        rails_oracle::CoutCapture cap;
        auto t0 = std::chrono::steady_clock::now();
        ret = solver.solve(V, T);
        rails_cuda_sync();
        if (seconds)
            *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        log = cap.str();
    }

    std::vector<double> ha, hr;
    int len = rails_oracle::parse_history(log, ha, hr);
    if (hist_len)
        *hist_len = len;
    for (int i = 0; i < std::min(len, hist_cap); ++i)
    {
        if (hist_abs) hist_abs[i] = ha[i];
        if (hist_rel) hist_rel[i] = hr[i];
    }

    int k = V.N();
    if (k_out)
        *k_out = k;
    if (V_out && T_out && k <= v_cap_cols)
    {
        V.to_host(V_out, m);
        for (int j = 0; j < k; ++j)
            for (int i = 0; i < k; ++i)
                T_out[i + (size_t)j * v_cap_cols] = T(i, j);
    }
    return ret;
}
