#include "CudaOperatorWrapper.hpp"
#include "OperatorEigs.hpp"

#include <cmath>
#include <iostream>
#include <vector>

namespace RAILS
{

namespace
{
void check(int rc, const char *what)
{
    if (rc)
        std::cerr << "CudaOperatorWrapper::" << what << " failed (" << rc << "): "
                  << rails_cuda_last_error() << std::endl;
}
}

CudaOperatorWrapper::CudaOperatorWrapper()
    : m_(-1), n_(-1), transpose_(false)
{}

CudaOperatorWrapper::CudaOperatorWrapper(int rows, int global_rows, int global_cols,
                                         int row_offset, const int *rowptr, const int *colidx,
                                         const double *vals)
    : m_(global_rows), n_(global_cols), transpose_(false)
{
    rails_csr_t h = nullptr;
    int rc = rails_csr_create(rows, global_rows, global_cols, row_offset, rowptr, colidx, vals, &h);
    check(rc, "constructor");
    if (!rc)
        ptr_ = std::make_shared<Handle>(h);
}

CudaOperatorWrapper::CudaOperatorWrapper(int m, int n, const int *rowptr, const int *colidx,
                                         const double *vals)
    : CudaOperatorWrapper(m, m, n, 0, rowptr, colidx, vals)
{}

CudaOperatorWrapper::CudaOperatorWrapper(int m, int n)
    : m_(m), n_(n), transpose_(false)
{
    std::vector<int> rowptr(m + 1), colidx((size_t)m * n);
    std::vector<double> vals((size_t)m * n, 0.0);
    for (int i = 0; i <= m; ++i)
        rowptr[i] = i * n;
    for (int i = 0; i < m; ++i)
        for (int j = 0; j < n; ++j)
            colidx[(size_t)i * n + j] = j;
    rails_csr_t h = nullptr;
    int rc = rails_csr_create(m, m, n, 0, rowptr.data(), colidx.data(), vals.data(), &h);
    check(rc, "constructor");
    if (!rc)
        ptr_ = std::make_shared<Handle>(h);
}

CudaOperatorWrapper::Element::operator double() const
{
    double v = 0.0;
    check(rails_csr_get_value(h_, row_, col_, &v), "operator()");
    return v;
}

CudaOperatorWrapper::Element &CudaOperatorWrapper::Element::operator=(double v)
{
    check(rails_csr_set_value(h_, row_, col_, v), "operator()");
    return *this;
}

CudaOperatorWrapper::Element CudaOperatorWrapper::operator()(int i, int j)
{
    return Element(ptr_ ? ptr_->h : nullptr, i, j);
}

double CudaOperatorWrapper::operator()(int i, int j) const
{
    return (double)Element(ptr_ ? ptr_->h : nullptr, i, j);
}

int CudaOperatorWrapper::eigs(CudaMultiVectorWrapper &V, HostDenseMatrix &D, int num, double tol) const
{
    return symmetric_eigs(*this, V, D, num, tol);
}

CudaOperatorWrapper::CudaOperatorWrapper(std::shared_ptr<const Action> const &action)
    : action_(action), m_(action ? action->rows() : -1), n_(action ? action->cols() : -1),
      transpose_(false)
{}

CudaOperatorWrapper CudaOperatorWrapper::transpose() const
{
    // flag on a shallow copy (src/Epetra_OperatorWrapper.cpp:50-55)
    CudaOperatorWrapper out(*this);
    out.transpose_ = !out.transpose_;
    return out;
}

CudaMultiVectorWrapper CudaOperatorWrapper::operator*(CudaMultiVectorWrapper const &other) const
{
    // result lives on the operator's range map with other.N() columns
    // (src/Epetra_OperatorWrapper.cpp:75-91)
    if (action_)
    {
        if (other.M() != N() || other.transpose_)
        {
            std::cerr << "Incomplatible sizes: operator " << M() << "x" << N() << " times "
                      << other.M() << "x" << other.N() << std::endl;
            return CudaMultiVectorWrapper(M(), other.N());
        }
        return action_->apply(other, transpose_);
    }
    other.materialize();
    CudaMultiVectorWrapper out;
    out.allocate(M(), other.N(), other.replicated_ && rails_cuda_comm_size() == 1);
    if (!ptr_ || other.M() != N() || other.transpose_)
    {
        std::cerr << "Incomplatible sizes: operator " << M() << "x" << N() << " times "
                  << other.M() << "x" << other.N() << std::endl;
        return out;
    }
    if (other.N() > 0 && out.buf_ && other.buf_)
        check(rails_csr_spmm(ptr_->h, transpose_ ? 1 : 0, other.buf_->h, other.col0_, other.n_,
                             out.buf_->h, 0), "operator*");
    return out;
}

double CudaOperatorWrapper::norm() const
{
    double out = 0.0;
    if (!ptr_)
    {
        if (action_)
            std::cerr << "CudaOperatorWrapper::norm is not available for a composite operator" << std::endl;
        return out;
    }
    if (n_ <= 512 && m_ <= 65536 && rails_cuda_comm_size() == 1)
    {
        // as the reference: A'A explicitly, all its eigenvalues, sqrt of the largest.
        // A'A comes from the same inner-product kernel MultiVector::dot uses.
        CudaMultiVectorWrapper dense(m_, n_);
        int rc = rails_csr_to_dense(ptr_->h, dense.device_handle(), dense.device_col0());
        check(rc, "norm");
        if (rc)
            return out;
        HostDenseMatrix gram = dense.dot(dense), vectors, values;
        gram.eigs(vectors, values);
        return std::sqrt(values.norm_inf());
    }
    check(rails_csr_norm2(ptr_->h, &out), "norm");
    return out;
}

int CudaOperatorWrapper::M() const { return transpose_ ? n_ : m_; }
int CudaOperatorWrapper::N() const { return transpose_ ? m_ : n_; }
long long CudaOperatorWrapper::nnz() const { return ptr_ ? rails_csr_nnz(ptr_->h) : 0; }

}
