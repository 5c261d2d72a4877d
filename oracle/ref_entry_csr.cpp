// TEST INFRASTRUCTURE (oracle/_ref build): the UNMODIFIED reference solver loop
// (/root/reference/src/LyapunovSolver.hpp) over the reference's own StlWrapper as
// MultiVector and DenseMatrix and a CSR + OpenMP Operator (oracle/cpu_backend.hpp) in
// place of the Epetra one, behind a C entry point.  This is the reference's CPU
// implementation of the path for sparse problems of any size:
//     Solver<CsrHostOperator, StlWrapper, StlWrapper>
// bench.py times it as `--impl reference` (cpu_baseline.kind "reference") wherever the
// reference loop can solve the workload (M = I: the C++ loop never reads M).
// Only the .so (oracle/_ref/librails_ref_csr.so) travels to the GPU box.
#include <algorithm>
#include <chrono>
#include <iostream>
#include <streambuf>
#include <utility>

#include "capture.hpp"
#include "cpu_backend.hpp"

#include "src/LyapunovSolver.hpp"
#include "src/StlWrapper.hpp"

using namespace RAILS;

namespace
{

// std::cout capture that also records WHEN each line ended: the reference reports
// progress only on std::cout (src/LyapunovSolver.hpp:219-221), one line per iteration,
// so the time stamps of the "Iteration" lines are the per-iteration clock
class TimedCapture : public std::streambuf
{
    std::string text_;
    std::vector<double> line_end_;
    std::streambuf *old_;
    std::streamsize old_prec_;
    std::chrono::steady_clock::time_point t0_;

protected:
    int overflow(int c) override
    {
        if (c != EOF)
        {
            text_.push_back((char)c);
            if (c == '\n')
                line_end_.push_back(std::chrono::duration<double>(std::chrono::steady_clock::now() - t0_).count());
        }
        return c;
    }

public:
    TimedCapture() : t0_(std::chrono::steady_clock::now())
    {
        old_ = std::cout.rdbuf(this);
        old_prec_ = std::cout.precision(17);
    }
    ~TimedCapture()
    {
        std::cout.rdbuf(old_);
        std::cout.precision(old_prec_);
    }
    // seconds since construction at which the i-th "Iteration " line was complete
    std::vector<double> iteration_times() const
    {
        std::vector<double> out;
        size_t pos = 0, line = 0;
        while (pos < text_.size() && line < line_end_.size())
        {
            size_t end = text_.find('\n', pos);
            if (end == std::string::npos) break;
            if (text_.compare(pos, 10, "Iteration ") == 0)
                out.push_back(line_end_[line]);
            pos = end + 1;
            ++line;
        }
        return out;
    }
    std::string const &str() const { return text_; }
};

}

extern "C" int ref_csr_solve(
    int m, const int *rowptr, const int *colidx, const double *vals,
    int n_b, const double *B_in, unsigned seed, int skip_draws,
    int nparams, const char *const *keys, const double *pvals,
    int hist_cap, double *hist_abs, double *hist_rel, double *hist_sec, int *hist_len,
    int *k_out, int v_cap_cols, double *V_out, double *T_out,
    double *seconds)
{
    std::srand(seed);
    for (int i = 0; i < skip_draws; ++i)
        (void)std::rand();      // the caller made B with `srand(seed); B.random()` itself

    rails_oracle::CsrHostOperator A(m, m, rowptr, colidx, vals);

    StlWrapper B(m, n_b);
    if (B_in)
    {
        double *b = B;
        for (int j = 0; j < n_b; ++j)
            std::copy(B_in + (size_t)j * m, B_in + (size_t)(j + 1) * m, b + (size_t)j * B.LDA());
    }
    else
        B.random();

    // M is stored and never read by the C++ solver
    Solver<rails_oracle::CsrHostOperator, StlWrapper, StlWrapper> solver(A, B, A);
    rails_oracle::ParameterList params(nparams, keys, pvals);
    int ret = solver.set_parameters(params);
    if (ret)
        return 100 + ret;

    StlWrapper V(m, 1);
    StlWrapper T;

    std::string log;
    std::vector<double> when;
    {
        TimedCapture cap;
        auto t0 = std::chrono::steady_clock::now();
        ret = solver.solve(V, T);
        auto t1 = std::chrono::steady_clock::now();
        if (seconds)
            *seconds = std::chrono::duration<double>(t1 - t0).count();
        log = cap.str();
        when = cap.iteration_times();
    }

    std::vector<double> ha, hr;
    int len = rails_oracle::parse_history(log, ha, hr);
    if (hist_len)
        *hist_len = len;
    for (int i = 0; i < std::min(len, hist_cap); ++i)
    {
        if (hist_abs) hist_abs[i] = ha[i];
        if (hist_rel) hist_rel[i] = hr[i];
        if (hist_sec) hist_sec[i] = i < (int)when.size() ? when[i] : 0.0;
    }

    int k = V.N();
    if (k_out)
        *k_out = k;
    if (V_out && T_out && k <= v_cap_cols)
    {
        const double *v = V;
        for (int j = 0; j < k; ++j)
            std::copy(v + (size_t)j * V.LDA(), v + (size_t)j * V.LDA() + m, V_out + (size_t)j * m);
        for (int j = 0; j < k; ++j)
            for (int i = 0; i < k; ++i)
                T_out[i + (size_t)j * v_cap_cols] = T(i, j);
    }
    return ret;
}
