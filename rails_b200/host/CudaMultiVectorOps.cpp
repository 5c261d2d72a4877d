// Arithmetic members of CudaMultiVectorWrapper: each is one or two calls into
// the C ABI (include/rails_cuda.h).  Reference counterparts are cited per
// member (StlWrapper = /root/reference/src/StlWrapper.cpp).
#include "CudaMultiVectorWrapper.hpp"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <iostream>

namespace RAILS
{

namespace
{

void check(int rc, const char *what)
{
    if (rc)
        std::cerr << "CudaMultiVectorWrapper::" << what << " failed (" << rc << "): "
                  << rails_cuda_last_error() << std::endl;
}

}

// ---- lazy products ---------------------------------------------------------------

void CudaMultiVectorWrapper::Lazy::evaluate()
{
    if (result)
        return;
    rails_mv_t h = nullptr;
    int rc = rails_mv_alloc(rails_mv_rows(src->h), n,
                            replicated ? RAILS_MV_REPLICATED : RAILS_MV_DISTRIBUTED, &h);
    check(rc, "lazy allocate");
    if (rc)
        return;
    result = std::make_shared<Buffer>(h);
    if (n == 0)
        return;
    if (!has_coef)
        rc = rails_mv_scaled_copy(h, 0, alpha, src->h, src_col0, n);
    else if (dev_coef)
        rc = rails_mv_gemm_dev(src->h, src_col0, src_k, dev_coef->h, dev_coef_col0, n, alpha, 0.0, h, 0);
    else
        rc = rails_mv_gemm(src->h, src_col0, src_k, coef, coef.LDA(), n, alpha, 0.0, h, 0);
    check(rc, "lazy evaluate");
}

bool CudaMultiVectorWrapper::Lazy::reads(Buffer const *b, int col0, int ncols) const
{
    if (src.get() == b)
    {
        int w = has_coef ? src_k : n;
        if (col0 < src_col0 + w && src_col0 < col0 + ncols)
            return true;
    }
    if (dev_coef.get() == b && col0 < dev_coef_col0 + n && dev_coef_col0 < col0 + ncols)
        return true;
    return false;
}

// Columns [col0, col0+ncols) are about to change: evaluate every pending
// product that still reads them.
void CudaMultiVectorWrapper::Buffer::before_write(int col0, int ncols)
{
    if (readers.empty())
        return;
    std::vector<std::weak_ptr<Lazy> > keep;
    std::vector<std::shared_ptr<Lazy> > run;
    for (auto &w : readers)
        if (auto l = w.lock())
        {
            if (l->result)
                continue;
            if (l->reads(this, col0, ncols))
                run.push_back(l);
            else
                keep.push_back(w);
        }
    readers.swap(keep);
    for (auto &l : run)
        l->evaluate();
}

void CudaMultiVectorWrapper::materialize() const
{
    if (!lazy_)
        return;
    lazy_->evaluate();
    buf_ = lazy_->result;
    lazy_.reset();
}

void CudaMultiVectorWrapper::will_write()
{
    materialize();
    if (buf_ && n_ > 0)
        buf_->before_write(col0_, n_);
}

CudaMultiVectorWrapper CudaMultiVectorWrapper::make_lazy(std::shared_ptr<Lazy> const &l)
{
    CudaMultiVectorWrapper out;
    out.lazy_ = l;
    out.col0_ = 0;
    out.m_ = l->m;
    out.n_ = l->n;
    out.n_max_ = l->n;
    out.replicated_ = l->replicated;
    l->src->readers.push_back(l);
    if (l->dev_coef)
        l->dev_coef->readers.push_back(l);
    return out;
}

// ---- BLAS-1 ----------------------------------------------------------------------

CudaMultiVectorWrapper &CudaMultiVectorWrapper::operator*=(double other)
{
    // DSCAL (:131-137)
    will_write();
    if (buf_ && n_ > 0)
        check(rails_mv_scal(buf_->h, col0_, n_, other), "*=");
    orthogonalized_ = 0;
    return *this;
}

CudaMultiVectorWrapper &CudaMultiVectorWrapper::operator/=(double other)
{
    // `*this *= 1.0 / other` (:139-143): the reciprocal is rounded first
    return *this *= 1.0 / other;
}

void CudaMultiVectorWrapper::add_scaled(CudaMultiVectorWrapper const &other, double sign)
{
    will_write();
    orthogonalized_ = 0;
    if (!buf_ || n_ <= 0)
        return;
    if (other.lazy_ && !other.lazy_->result)
    {
        Lazy const &l = *other.lazy_;
        bool alias = l.reads(buf_.get(), col0_, n_);
        if (!alias && l.has_coef)
        {
            // y += V * C as one pass over V: sum first, then one rounded add
            int rc = l.dev_coef
                ? rails_mv_gemm_dev(l.src->h, l.src_col0, l.src_k, l.dev_coef->h, l.dev_coef_col0,
                                    n_, sign * l.alpha, 1.0, buf_->h, col0_)
                : rails_mv_gemm(l.src->h, l.src_col0, l.src_k, l.coef, l.coef.LDA(), n_,
                                sign * l.alpha, 1.0, buf_->h, col0_);
            check(rc, "+= V * C");
            return;
        }
        if (!alias)
        {
            check(rails_mv_add_scaled(buf_->h, col0_, sign, l.alpha, l.src->h, l.src_col0, n_),
                  "+= alpha * x");
            return;
        }
    }
    other.materialize();
    // DAXPY with +-1 (:147-161)
    check(rails_mv_axpy(buf_->h, col0_, sign, other.buf_->h, other.col0_, n_), "+=");
}

CudaMultiVectorWrapper &CudaMultiVectorWrapper::operator+=(CudaMultiVectorWrapper const &other)
{
    add_scaled(other, 1.0);
    return *this;
}

CudaMultiVectorWrapper &CudaMultiVectorWrapper::operator-=(CudaMultiVectorWrapper const &other)
{
    add_scaled(other, -1.0);
    return *this;
}

CudaMultiVectorWrapper CudaMultiVectorWrapper::operator+(CudaMultiVectorWrapper const &other) const
{
    CudaMultiVectorWrapper out(*this);
    out += other;
    return out;
}

CudaMultiVectorWrapper operator*(double d, CudaMultiVectorWrapper const &other)
{
    // scaled copy (:481-487), recorded lazily so `y -= d * x` is one kernel
    other.materialize();
    auto l = std::make_shared<CudaMultiVectorWrapper::Lazy>();
    l->src = other.buf_;
    l->src_col0 = other.col0_;
    l->src_k = other.n_;
    l->has_coef = false;
    l->alpha = d;
    l->m = other.m_;
    l->n = other.n_;
    l->replicated = other.replicated_;
    CudaMultiVectorWrapper out = CudaMultiVectorWrapper::make_lazy(l);
    out.transpose_ = other.transpose_;
    return out;
}

// ---- products ---------------------------------------------------------------------

CudaMultiVectorWrapper CudaMultiVectorWrapper::operator*(HostDenseMatrix const &other) const
{
    // V * C (:168-187; src/Epetra_MultiVectorWrapper.cpp:194-219)
    materialize();
    if (other.M() != N() || transpose_)
    {
        std::cerr << "Incomplatible matrices of sizes " << M() << "x" << N() << " and "
                  << other.M() << "x" << other.N() << std::endl;
        return CudaMultiVectorWrapper(*this, other.N());
    }
    auto l = std::make_shared<Lazy>();
    l->src = buf_;
    l->src_col0 = col0_;
    l->src_k = n_;
    l->has_coef = true;
    // private compact copy: the caller may reuse its matrix (Z = T * Z)
    if (other.transposed())
    {
        HostDenseMatrix c(other.M(), other.N());
        HostDenseMatrix plain = other.transpose();
        for (int j = 0; j < other.N(); ++j)
            for (int i = 0; i < other.M(); ++i)
                c(i, j) = plain(j, i);
        l->coef = c;
    }
    else
        l->coef = other.copy();
    l->alpha = 1.0;
    l->m = m_;
    l->n = other.N();
    l->replicated = replicated_;
    return make_lazy(l);
}

CudaMultiVectorWrapper CudaMultiVectorWrapper::operator*(CudaMultiVectorWrapper const &other) const
{
    materialize();
    other.materialize();
    if (other.M() != N() || other.transpose_)
    {
        std::cerr << "Incomplatible matrices of sizes " << M() << "x" << N() << " and "
                  << other.M() << "x" << other.N()
                  << (other.transpose_ ? " (right operand transposed: m x m result not formed on the device)" : "")
                  << std::endl;
        return CudaMultiVectorWrapper(*this, other.N());
    }
    if (transpose_)
    {
        // B' * W: n x e block, identical on every rank (Epetra LocalMap result,
        // src/Epetra_MultiVectorWrapper.cpp:221-241); stays on the device
        CudaMultiVectorWrapper out = replicated(n_, other.n_);
        if (out.buf_ && n_ > 0 && other.n_ > 0)
            check(rails_mv_dot_dev(buf_->h, col0_, n_, other.buf_->h, other.col0_, other.n_,
                                   out.buf_->h, 0), "B' * W");
        return out;
    }
    // B * Y with Y a small replicated block (src/LyapunovSolver.hpp:390)
    auto l = std::make_shared<Lazy>();
    l->src = buf_;
    l->src_col0 = col0_;
    l->src_k = n_;
    l->has_coef = true;
    l->dev_coef = other.buf_;
    l->dev_coef_col0 = other.col0_;
    l->alpha = 1.0;
    l->m = m_;
    l->n = other.n_;
    l->replicated = replicated_;
    return make_lazy(l);
}

HostDenseMatrix CudaMultiVectorWrapper::dot(CudaMultiVectorWrapper const &other) const
{
    // X' Y -> host (:394-412)
    materialize();
    other.materialize();
    HostDenseMatrix out(N(), other.N());
    if (other.M() != M())
    {
        std::cerr << "Incomplatible matrices of sizes " << M() << "x" << N() << " and "
                  << other.M() << "x" << other.N() << std::endl;
        return out;
    }
    if (N() > 0 && other.N() > 0 && buf_ && other.buf_)
        check(rails_mv_dot(buf_->h, col0_, n_, other.buf_->h, other.col0_, other.n_, out, out.LDA()),
              "dot");
    return out;
}

double CudaMultiVectorWrapper::norm() const
{
    // 2-norm: sqrt of the largest eigenvalue of X'X (:265-289); for a single
    // column this is sqrt(|x'x|)
    HostDenseMatrix g = dot(*this);
    if (n_ == 1)
        return std::sqrt(std::abs(g(0, 0)));
    HostDenseMatrix v, d;
    g.eigs(v, d);
    double mx = 0.0;
    for (int i = 0; i < n_; ++i)
        mx = std::max(mx, std::sqrt(std::abs(d(i, 0))));
    return mx;
}

double CudaMultiVectorWrapper::norm_inf() const
{
    materialize();
    double out = 0.0;
    if (buf_ && n_ > 0)
        check(rails_mv_norm_inf(buf_->h, col0_, n_, &out), "norm_inf");
    return out;
}

void CudaMultiVectorWrapper::orthogonalize()
{
    // column-wise CGS2 from the first non-orthogonal column (:305-321), as
    // one device-side chain
    will_write();
    if (buf_ && orthogonalized_ < N())
        check(rails_mv_orthogonalize(buf_->h, col0_, orthogonalized_, N() - 1), "orthogonalize");
    orthogonalized_ = N();
}

void CudaMultiVectorWrapper::random()
{
    // one std::rand() per call seeds the engine (:414-423)
    will_write();
    unsigned int seed = (unsigned int)std::rand();
    if (buf_ && n_ > 0)
    {
        long long off = 0, end = 0;
        if (!replicated_ && rails_cuda_comm_size() > 1)
            rails_partition_rows(m_, rails_cuda_comm_size(), rails_cuda_comm_rank(), &off, &end);
        check(rails_mv_random(buf_->h, col0_, n_, off, seed), "random");
    }
    orthogonalized_ = 0;
}

}
