// Operator concept of RAILS::Solver for the B200 backend: a CSR matrix in HBM.
//
// Mirrors Epetra_OperatorWrapper (/root/reference/src/Epetra_OperatorWrapper.hpp,
// .cpp:41-145): shallow handle copy, transpose() flips a flag on a copy,
// operator*(MultiVector) applies the operator, norm() is the 2-norm, M()/N()
// swap under transpose.  The reference converts implicitly from its native
// handle (Teuchos::RCP<Epetra_Operator>); here the native input is a host CSR
// triple (rowptr, colidx, vals), uploaded once at construction.
#ifndef RAILS_B200_CUDAOPERATORWRAPPER_HPP
#define RAILS_B200_CUDAOPERATORWRAPPER_HPP

#include "CudaMultiVectorWrapper.hpp"
#include "rails_cuda.h"

#include <memory>

namespace RAILS
{

class CudaOperatorWrapper
{
public:
    // An operator that is not one CSR matrix: the Schur complement
    // A22 - A21 A11^-1 A12 and its inverse (CudaSchurOperator.hpp).  The
    // reference wraps any Epetra_Operator the same way -- its SchurOperator is
    // handed to the solver through Epetra_OperatorWrapper (src/main.cpp:90-98).
    struct Action
    {
        virtual ~Action() {}
        virtual CudaMultiVectorWrapper apply(CudaMultiVectorWrapper const &x, bool transpose) const = 0;
        virtual int rows() const = 0;
        virtual int cols() const = 0;
    };

private:
    struct Handle
    {
        rails_csr_t h;
        explicit Handle(rails_csr_t handle) : h(handle) {}
        ~Handle() { rails_csr_free(h); }
        Handle(Handle const &) = delete;
        Handle &operator=(Handle const &) = delete;
    };
    std::shared_ptr<Handle> ptr_;
    std::shared_ptr<const Action> action_;
    int m_, n_;          // GLOBAL sizes
    bool transpose_;

public:
    CudaOperatorWrapper();
    // this rank's rows [row_offset, row_offset + rows) of a global_rows x
    // global_cols matrix; column indices are global.  With one rank pass
    // rows = global_rows, row_offset = 0.
    CudaOperatorWrapper(int rows, int global_rows, int global_cols, int row_offset,
                        const int *rowptr, const int *colidx, const double *vals);
    // convenience: whole matrix on one rank
    CudaOperatorWrapper(int m, int n, const int *rowptr, const int *colidx, const double *vals);
    // composite operator (implicit, like the reference's conversion from
    // Teuchos::RCP<Epetra_Operator>, src/Epetra_OperatorWrapper.hpp)
    CudaOperatorWrapper(std::shared_ptr<const Action> const &action);
    bool valid() const { return ptr_ || action_; }

    virtual ~CudaOperatorWrapper() {}

    CudaOperatorWrapper transpose() const;
    CudaMultiVectorWrapper operator*(CudaMultiVectorWrapper const &other) const;

    double norm() const;

    int M() const;
    int N() const;

    long long nnz() const;
    rails_csr_t device_handle() const { return ptr_ ? ptr_->h : nullptr; }
};

}

#endif
