#include "CudaMultiVectorWrapper.hpp"

#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdlib>
#include <iostream>
#include <vector>

namespace RAILS
{

namespace
{

// Reference style: report and carry on (src/StlWrapper.cpp:173-179)
void check(int rc, const char *what)
{
    if (rc)
        std::cerr << "CudaMultiVectorWrapper::" << what << " failed (" << rc << "): "
                  << rails_cuda_last_error() << std::endl;
}

long long local_rows_of(int m, bool replicated)
{
    if (replicated || rails_cuda_comm_size() == 1)
        return m;
    long long b = 0, e = 0;
    rails_partition_rows(m, rails_cuda_comm_size(), rails_cuda_comm_rank(), &b, &e);
    return e - b;
}

long long row_offset_of(int m, bool replicated)
{
    if (replicated || rails_cuda_comm_size() == 1)
        return 0;
    long long b = 0, e = 0;
    rails_partition_rows(m, rails_cuda_comm_size(), rails_cuda_comm_rank(), &b, &e);
    return b;
}

}

// ---- test-only element proxy ------------------------------------------------

CudaMultiVectorWrapper::Element::operator double() const
{
    double v = 0.0;
    check(rails_mv_get(h_, row_, col_, &v), "operator()");
    return v;
}

CudaMultiVectorWrapper::Element &CudaMultiVectorWrapper::Element::operator=(double v)
{
    check(rails_mv_set(h_, row_, col_, v), "operator()");
    return *this;
}

// ---- construction -------------------------------------------------------------

CudaMultiVectorWrapper::CudaMultiVectorWrapper()
    : col0_(0), m_(-1), n_(-1), n_max_(-1), orthogonalized_(0), is_view_(false),
      transpose_(false), replicated_(false)
{}

void CudaMultiVectorWrapper::allocate(int m, int n, bool replicated)
{
    rails_mv_t h = nullptr;
    int rc = rails_mv_alloc(local_rows_of(m, replicated), n,
                            replicated ? RAILS_MV_REPLICATED : RAILS_MV_DISTRIBUTED, &h);
    check(rc, "allocate");
    buf_ = rc ? nullptr : std::make_shared<Buffer>(h);
    col0_ = 0;
    m_ = m;
    n_ = n;
    n_max_ = n;
    replicated_ = replicated;
    lazy_.reset();
}

CudaMultiVectorWrapper::CudaMultiVectorWrapper(int m, int n)
    : CudaMultiVectorWrapper()
{
    allocate(m, n, false);
}

CudaMultiVectorWrapper CudaMultiVectorWrapper::replicated(int m, int n)
{
    CudaMultiVectorWrapper out;
    out.allocate(m, n, true);
    return out;
}

CudaMultiVectorWrapper CudaMultiVectorWrapper::from_host(int m, int n, const double *host,
                                                         long long ldh)
{
    CudaMultiVectorWrapper out(m, n);
    if (out.buf_ && n > 0)
        check(rails_mv_upload(out.buf_->h, 0, n, host + row_offset_of(m, false), ldh), "from_host");
    return out;
}

// deep copy (src/StlWrapper.cpp:31-44); a copy of a view owns just its columns
CudaMultiVectorWrapper::CudaMultiVectorWrapper(CudaMultiVectorWrapper const &other)
    : CudaMultiVectorWrapper()
{
    other.materialize();
    m_ = other.m_;
    n_ = other.n_;
    n_max_ = other.n_;
    orthogonalized_ = other.orthogonalized_;
    transpose_ = other.transpose_;
    replicated_ = other.replicated_;
    if (other.buf_)
    {
        allocate(other.m_, std::max(other.n_, 0), other.replicated_);
        orthogonalized_ = other.orthogonalized_;
        transpose_ = other.transpose_;
        if (buf_ && n_ > 0)
            check(rails_mv_copy_cols(buf_->h, 0, other.buf_->h, other.col0_, n_), "copy constructor");
    }
}

CudaMultiVectorWrapper::CudaMultiVectorWrapper(CudaMultiVectorWrapper &&o) noexcept
    : buf_(std::move(o.buf_)), col0_(o.col0_), m_(o.m_), n_(o.n_), n_max_(o.n_max_),
      orthogonalized_(o.orthogonalized_), is_view_(o.is_view_), transpose_(o.transpose_),
      replicated_(o.replicated_), lazy_(std::move(o.lazy_))
{}

CudaMultiVectorWrapper::CudaMultiVectorWrapper(CudaMultiVectorWrapper const &other, int n)
    : CudaMultiVectorWrapper()
{
    // same rows (and row distribution) as `other`, n fresh columns (:46-51)
    allocate(other.M(), n, other.transpose_ ? true : other.replicated_);
}

rails_mv_t CudaMultiVectorWrapper::handle() const
{
    materialize();
    return buf_ ? buf_->h : nullptr;
}

// ---- assignment ---------------------------------------------------------------

void CudaMultiVectorWrapper::assign_columns(CudaMultiVectorWrapper const &other)
{
    // `view = expr`: copy into place (src/StlWrapper.cpp:80-92)
    assert(N() == other.N());
    will_write();
    if (other.lazy_ && !other.lazy_->result && other.lazy_->has_coef && buf_)
    {
        // view = V * C  ->  one kernel straight into the destination, unless
        // the destination overlaps the columns being read
        Lazy const &l = *other.lazy_;
        bool alias = l.reads(buf_.get(), col0_, n_);
        if (!alias)
        {
            int rc;
            if (l.dev_coef)
                rc = rails_mv_gemm_dev(l.src->h, l.src_col0, l.src_k, l.dev_coef->h, l.dev_coef_col0,
                                       n_, l.alpha, 0.0, buf_->h, col0_);
            else
                rc = rails_mv_gemm(l.src->h, l.src_col0, l.src_k, l.coef, l.coef.LDA(), n_, l.alpha,
                                   0.0, buf_->h, col0_);
            check(rc, "view = V * C");
            return;
        }
    }
    other.materialize();
    int n = std::min(n_, other.n_);
    if (n > 0 && buf_ && other.buf_)
        check(rails_mv_copy_cols(buf_->h, col0_, other.buf_->h, other.col0_, n), "view assignment");
}

CudaMultiVectorWrapper &CudaMultiVectorWrapper::operator=(CudaMultiVectorWrapper const &other)
{
    if (this == &other)
        return *this;
    if (!is_view_)
    {
        // shallow: share the buffer (src/StlWrapper.cpp:68-78)
        buf_ = other.buf_;
        lazy_ = other.lazy_;
        col0_ = other.col0_;
        m_ = other.m_;
        n_ = other.n_;
        n_max_ = other.n_max_;
        orthogonalized_ = other.orthogonalized_;
        transpose_ = other.transpose_;
        replicated_ = other.replicated_;
        return *this;
    }
    assign_columns(other);
    return *this;
}

CudaMultiVectorWrapper &CudaMultiVectorWrapper::operator=(CudaMultiVectorWrapper &other)
{
    return *this = static_cast<CudaMultiVectorWrapper const &>(other);
}

CudaMultiVectorWrapper &CudaMultiVectorWrapper::operator=(CudaMultiVectorWrapper &&other)
{
    if (this == &other)
        return *this;
    if (!is_view_)
    {
        buf_ = std::move(other.buf_);
        lazy_ = std::move(other.lazy_);
        col0_ = other.col0_;
        m_ = other.m_;
        n_ = other.n_;
        n_max_ = other.n_max_;
        orthogonalized_ = other.orthogonalized_;
        transpose_ = other.transpose_;
        replicated_ = other.replicated_;
        return *this;
    }
    assign_columns(other);
    return *this;
}

CudaMultiVectorWrapper &CudaMultiVectorWrapper::operator=(double other)
{
    will_write();
    if (buf_ && n_ > 0)
        check(rails_mv_fill(buf_->h, col0_, n_, other), "= double");
    orthogonalized_ = 0;
    return *this;
}

// ---- sizes, views ------------------------------------------------------------

int CudaMultiVectorWrapper::M() const { return transpose_ ? n_ : m_; }
int CudaMultiVectorWrapper::N() const { return transpose_ ? m_ : n_; }

long long CudaMultiVectorWrapper::local_rows() const
{
    return buf_ ? rails_mv_rows(buf_->h) : (lazy_ && lazy_->src ? rails_mv_rows(lazy_->src->h) : 0);
}

void CudaMultiVectorWrapper::resize(int n)
{
    resize(m_, n);
}

void CudaMultiVectorWrapper::resize(int m, int n)
{
    materialize();
    orthogonalized_ = std::min(orthogonalized_, n);
    if (buf_ && m == m_ && n <= n_max_)
    {
        n_ = n;
        return;
    }
    // grow the column capacity: new allocation, old columns carried over
    // (src/StlWrapper.cpp:236-246); a different row count starts afresh with
    // the reference's warning (:249-262)
    std::shared_ptr<Buffer> old = buf_;
    int old_col0 = col0_, old_n = std::max(n_, 0), old_m = m_;
    bool repl = replicated_;
    int orth = orthogonalized_;
    bool tr = transpose_;
    allocate(m, n, repl);
    orthogonalized_ = orth;
    transpose_ = tr;
    is_view_ = false;
    if (old && old_m == m)
    {
        int ncopy = std::min(old_n, n);
        if (ncopy > 0 && buf_)
            check(rails_mv_copy_cols(buf_->h, 0, old->h, old_col0, ncopy), "resize");
    }
    else if (old && old_m > 0)
        std::cerr << "Warning: data dropped during resize from " << old_m << " to " << m
                  << " rows" << std::endl;
}

CudaMultiVectorWrapper CudaMultiVectorWrapper::view(int m, int n)
{
    return static_cast<CudaMultiVectorWrapper const *>(this)->view(m, n);
}

CudaMultiVectorWrapper CudaMultiVectorWrapper::view(int m, int n) const
{
    // view() = all columns, view(i) = column i, view(a, b) = columns a..b
    // inclusive and only when b > 0 (src/StlWrapper.cpp:323-340)
    int num = 1;
    if (n > 0 && m >= 0)
        num = n - m + 1;
    else if (m < 0)
    {
        m = 0;
        num = N();
    }
    if (lazy_ && !lazy_->result && lazy_->has_coef && !lazy_->dev_coef && !transpose_ &&
        m + num <= n_)
    {
        // columns of a pending product V * C are the product with the matching
        // columns of C: `eigenvectors.view(i)` (src/LyapunovSolver.hpp:339)
        // then costs one m x 1 product instead of the whole m x L block
        auto l = std::make_shared<Lazy>(*lazy_);
        HostDenseMatrix c(lazy_->coef.M(), num);
        for (int j = 0; j < num; ++j)
            for (int i = 0; i < lazy_->coef.M(); ++i)
                c(i, j) = lazy_->coef(i, m + j);
        l->coef = c;
        l->n = num;
        l->result.reset();
        CudaMultiVectorWrapper out = make_lazy(l);
        out.orthogonalized_ = 0;
        return out;
    }
    materialize();
    CudaMultiVectorWrapper out;
    out.buf_ = buf_;
    out.col0_ = col0_ + m;
    out.m_ = m_;
    out.n_ = num;
    out.n_max_ = num;
    out.orthogonalized_ = orthogonalized_;
    out.transpose_ = transpose_;
    out.replicated_ = replicated_;
    out.is_view_ = true;
    return out;
}

CudaMultiVectorWrapper CudaMultiVectorWrapper::copy() const
{
    return *this;
}

CudaMultiVectorWrapper CudaMultiVectorWrapper::transpose() const
{
    // a flag on a handle that shares the storage; the reference deep-copies
    // here (src/StlWrapper.cpp:425-431), which only costs time
    materialize();
    CudaMultiVectorWrapper out;
    out.buf_ = buf_;
    out.col0_ = col0_;
    out.m_ = m_;
    out.n_ = n_;
    out.n_max_ = n_max_;
    out.orthogonalized_ = orthogonalized_;
    out.replicated_ = replicated_;
    out.transpose_ = !transpose_;
    return out;
}

void CudaMultiVectorWrapper::push_back(CudaMultiVectorWrapper const &other)
{
    int n = N();
    int other_n = other.N();
    if (other.lazy_ && !other.lazy_->result && other.lazy_->has_coef && !other.lazy_->dev_coef &&
        other_n > 0)
    {
        // push_back(V * C): the product is formed directly in the new columns
        std::shared_ptr<Lazy> l = other.lazy_;
        resize(other_n + n);
        if (buf_ && !l->reads(buf_.get(), col0_ + n, other_n))
        {
            buf_->before_write(col0_ + n, other_n);
            check(rails_mv_gemm(l->src->h, l->src_col0, l->src_k, l->coef, l->coef.LDA(), other_n,
                                l->alpha, 0.0, buf_->h, col0_ + n), "push_back(V * C)");
            return;
        }
        n_ = n;                 // fall through to the general path
    }
    other.materialize();
    resize(other_n + n);
    if (other_n > 0 && buf_)
        buf_->before_write(col0_ + n, other_n);
    if (other_n > 0 && buf_ && other.buf_)
        check(rails_mv_copy_cols(buf_->h, col0_ + n, other.buf_->h, other.col0_, other_n), "push_back");
}

int CudaMultiVectorWrapper::to_host(double *host, long long ldh) const
{
    materialize();
    if (!buf_ || n_ <= 0)
        return 0;
    int rc = rails_mv_download(buf_->h, col0_, n_, host, ldh);
    check(rc, "to_host");
    return rc;
}

CudaMultiVectorWrapper::Element CudaMultiVectorWrapper::operator()(int i, int j)
{
    will_write();
    orthogonalized_ = 0;
    return Element(buf_->h, i, col0_ + j);
}

double CudaMultiVectorWrapper::operator()(int i, int j) const
{
    materialize();
    return (double)Element(buf_->h, i, col0_ + j);
}

}
