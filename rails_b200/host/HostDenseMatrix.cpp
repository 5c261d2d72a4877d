#include "HostDenseMatrix.hpp"

#include "lapack_decl.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <iostream>
#include <utility>

namespace RAILS
{

HostDenseMatrix::HostDenseMatrix()
    : m_(-1), n_(-1), ld_(-1), n_cap_(-1), is_view_(false), transpose_(false)
{}

HostDenseMatrix::HostDenseMatrix(int m, int n)
    : buf_(std::make_shared<std::vector<double> >((size_t)std::max(m, 1) * std::max(n, 1), 0.0)),
      m_(m), n_(n), ld_(std::max(m, 1)), n_cap_(n), is_view_(false), transpose_(false)
{}

HostDenseMatrix::HostDenseMatrix(HostDenseMatrix const &o)
    : m_(o.m_), n_(o.n_), ld_(o.ld_), n_cap_(o.n_cap_), is_view_(false), transpose_(o.transpose_)
{
    if (o.buf_)
        buf_ = std::make_shared<std::vector<double> >(*o.buf_);
}

HostDenseMatrix::HostDenseMatrix(HostDenseMatrix &&o) noexcept
    : buf_(std::move(o.buf_)), m_(o.m_), n_(o.n_), ld_(o.ld_), n_cap_(o.n_cap_),
      is_view_(o.is_view_), transpose_(o.transpose_)
{}

HostDenseMatrix &HostDenseMatrix::operator=(HostDenseMatrix const &o)
{
    if (this == &o)
        return *this;
    if (!is_view_)
    {
        buf_ = o.buf_;
        m_ = o.m_; n_ = o.n_; ld_ = o.ld_; n_cap_ = o.n_cap_;
        transpose_ = o.transpose_;
        return *this;
    }
    // view: copy the entries in place
    for (int j = 0; j < std::min(n_, o.n_); ++j)
        for (int i = 0; i < std::min(m_, o.m_); ++i)
            (*buf_)[i + (size_t)j * ld_] = (*o.buf_)[i + (size_t)j * o.ld_];
    return *this;
}

HostDenseMatrix &HostDenseMatrix::operator=(HostDenseMatrix &&o) noexcept
{
    if (this == &o)
        return *this;
    if (!is_view_)
    {
        buf_ = std::move(o.buf_);
        m_ = o.m_; n_ = o.n_; ld_ = o.ld_; n_cap_ = o.n_cap_;
        transpose_ = o.transpose_;
        return *this;
    }
    return *this = static_cast<HostDenseMatrix const &>(o);
}

HostDenseMatrix &HostDenseMatrix::operator=(double value)
{
    if (buf_)
        for (int j = 0; j < n_; ++j)
            std::fill_n(buf_->data() + (size_t)j * ld_, ld_, value);
    return *this;
}

HostDenseMatrix &HostDenseMatrix::operator*=(double value)
{
    if (buf_)
        for (int j = 0; j < n_; ++j)
            for (int i = 0; i < ld_; ++i)
                (*buf_)[i + (size_t)j * ld_] *= value;
    return *this;
}

HostDenseMatrix &HostDenseMatrix::operator/=(double value)
{
    return *this *= 1.0 / value;
}

HostDenseMatrix HostDenseMatrix::operator*(HostDenseMatrix const &o) const
{
    HostDenseMatrix out(M(), o.N());
    if (o.M() != N())
    {
        std::cerr << "Incomplatible matrices of sizes " << M() << "x" << N() << " and "
                  << o.M() << "x" << o.N() << std::endl;
        return out;
    }
    if (M() == 0 || o.N() == 0)
        return out;
    if (N() == 0)
    {
        out = 0.0;
        return out;
    }
    char ta = transpose_ ? 'T' : 'N', tb = o.transpose_ ? 'T' : 'N';
    int mm = M(), nn = o.N(), kk = N();
    int lda = ld_, ldb = o.ld_, ldc = out.ld_;
    if ((long long)mm * nn * kk <= 200000)
    {
        // k x k times k x 1 (Z = T * Z, twice per Lanczos step): a plain loop;
        // a threaded BLAS would wake its whole pool for ~100 flops
        const double *a = buf_->data(), *b = o.buf_->data();
        double *c = out.buf_->data();
        for (int j = 0; j < nn; ++j)
            for (int i = 0; i < mm; ++i)
            {
                double sum = 0.0;
                for (int l = 0; l < kk; ++l)
                    sum += (transpose_ ? a[l + (size_t)i * lda] : a[i + (size_t)l * lda]) *
                           (o.transpose_ ? b[j + (size_t)l * ldb] : b[l + (size_t)j * ldb]);
                c[i + (size_t)j * ldc] = sum;
            }
        return out;
    }
    double one = 1.0, zero = 0.0;
    dgemm_(&ta, &tb, &mm, &nn, &kk, &one, buf_->data(), &lda, o.buf_->data(), &ldb, &zero,
           out.buf_->data(), &ldc);
    return out;
}

double &HostDenseMatrix::operator()(int i, int j) { return (*buf_)[i + (size_t)j * ld_]; }
double const &HostDenseMatrix::operator()(int i, int j) const { return (*buf_)[i + (size_t)j * ld_]; }

HostDenseMatrix::operator double *() const { return buf_ ? buf_->data() : nullptr; }

void HostDenseMatrix::resize(int m, int n)
{
    if (buf_ && m <= ld_ && n <= n_cap_)
    {
        m_ = m; n_ = n;
        return;
    }
    // grow: entries keep their (i, j) position
    HostDenseMatrix out(m, n);
    if (buf_)
        for (int j = 0; j < std::min(n_, n); ++j)
            for (int i = 0; i < std::min(m_, m); ++i)
                out(i, j) = (*this)(i, j);
    buf_ = out.buf_;
    m_ = m; n_ = n; ld_ = out.ld_; n_cap_ = out.n_cap_;
}

double HostDenseMatrix::norm_inf() const
{
    double out = 0.0;
    for (int i = 0; i < m_; ++i)
    {
        double s = 0.0;
        for (int j = 0; j < n_; ++j)
            s += std::abs((*this)(i, j));
        out = std::max(out, s);
    }
    return out;
}

HostDenseMatrix HostDenseMatrix::view()
{
    HostDenseMatrix out;
    out.buf_ = buf_;
    out.m_ = m_; out.n_ = n_; out.ld_ = ld_; out.n_cap_ = n_cap_;
    out.transpose_ = transpose_;
    out.is_view_ = true;
    return out;
}

HostDenseMatrix HostDenseMatrix::copy() const
{
    // compact deep copy of the logical block
    HostDenseMatrix out(m_, n_);
    for (int j = 0; j < n_; ++j)
        for (int i = 0; i < m_; ++i)
            out(i, j) = (*this)(i, j);
    out.transpose_ = transpose_;
    return out;
}

HostDenseMatrix HostDenseMatrix::transpose() const
{
    HostDenseMatrix out;
    out.buf_ = buf_;
    out.m_ = m_; out.n_ = n_; out.ld_ = ld_; out.n_cap_ = n_cap_;
    out.transpose_ = !transpose_;
    return out;
}

int HostDenseMatrix::M() const { return transpose_ ? n_ : m_; }
int HostDenseMatrix::N() const { return transpose_ ? m_ : n_; }
int HostDenseMatrix::LDA() const { return ld_; }

int HostDenseMatrix::eigs(HostDenseMatrix &V, HostDenseMatrix &D, int num, double tol) const
{
    V = copy();
    int m = V.M();
    if (num < 1)
        num = m;
    D = HostDenseMatrix(m, 1);
    int info = 0;
    if (m > 0)
    {
        char jobz = 'V', uplo = 'U';
        int lda = V.LDA(), lwork = -1;
        double wq = 0.0;
        dsyev_(&jobz, &uplo, &m, V.buf_->data(), &lda, D.buf_->data(), &wq, &lwork, &info);
        lwork = std::max(3 * m, (int)wq);
        std::vector<double> work(lwork);
        dsyev_(&jobz, &uplo, &m, V.buf_->data(), &lda, D.buf_->data(), work.data(), &lwork, &info);
    }
    if (num != m || tol > 1e-14)
    {
        std::vector<std::pair<double, int> > order;
        for (int i = 0; i < m; ++i)
            order.push_back(std::make_pair(std::abs(D(i, 0)), i));
        std::stable_sort(order.begin(), order.end(),
                         [](std::pair<double, int> const &a, std::pair<double, int> const &b)
                         { return a.first > b.first; });
        HostDenseMatrix tv(m, num), td(num, 1);
        int idx = 0;
        for (int i = 0; i < std::min(num, m); ++i)
        {
            int src = order[i].second;
            if (std::abs(D(src, 0)) > tol)
            {
                for (int r = 0; r < m; ++r)
                    tv(r, idx) = V(r, src);
                td(idx, 0) = D(src, 0);
                idx++;
            }
        }
        tv.resize(m, idx);
        td.resize(idx, 1);
        V = tv;
        D = td;
    }
    if (info)
        std::cerr << "Eigenvalues info = " << info << std::endl;
    return info;
}

}
