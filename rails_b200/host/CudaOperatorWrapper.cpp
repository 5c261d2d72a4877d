#include "CudaOperatorWrapper.hpp"

#include <iostream>

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
    if (ptr_)
        check(rails_csr_norm2(ptr_->h, &out), "norm");
    else if (action_)
        std::cerr << "CudaOperatorWrapper::norm is not available for a composite operator" << std::endl;
    return out;
}

int CudaOperatorWrapper::M() const { return transpose_ ? n_ : m_; }
int CudaOperatorWrapper::N() const { return transpose_ ? m_ : n_; }
long long CudaOperatorWrapper::nnz() const { return ptr_ ? rails_csr_nnz(ptr_->h) : 0; }

}
