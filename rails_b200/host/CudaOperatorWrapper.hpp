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

    // test-only element access, entries of the sparsity pattern (the reference
    // keeps `(m, n)` + operator()(i, j) on its Operator for the typed tests,
    // src/Epetra_OperatorWrapper.cpp:224-282)
    class Element
    {
        rails_csr_t h_;
        int row_, col_;
    public:
        Element(rails_csr_t h, int row, int col) : h_(h), row_(row), col_(col) {}
        operator double() const;
        Element &operator=(double v);
        Element &operator=(Element const &o) { return *this = (double)o; }
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
    // test-only: an m x n operator whose sparsity pattern is FULL, all entries
    // zero, to be filled through operator()(i, j) (the reference inserts all n
    // columns of every row, src/Epetra_OperatorWrapper.cpp:224-250)
    CudaOperatorWrapper(int m, int n);
    Element operator()(int i, int j = 0);
    double operator()(int i, int j = 0) const;
    bool valid() const { return ptr_ || action_; }

    virtual ~CudaOperatorWrapper() {}

    CudaOperatorWrapper transpose() const;
    CudaMultiVectorWrapper operator*(CudaMultiVectorWrapper const &other) const;

    // 2-norm.  Small operators as the reference does it -- A'A formed explicitly,
    // all eigenvalues, sqrt of the largest (src/Epetra_OperatorWrapper.cpp:115-145)
    // -- larger ones by power iteration on A'A.
    double norm() const;

    // The `num` eigenvalues of largest magnitude of a SYMMETRIC operator and their
    // eigenvectors (src/Epetra_OperatorWrapper.cpp:147-222, Anasazi block
    // Krylov-Schur there; not used by the solver loop): OperatorEigs.hpp.
    int eigs(CudaMultiVectorWrapper &V, HostDenseMatrix &D, int num, double tol = 1e-6) const;

    int M() const;
    int N() const;

    long long nnz() const;
    rails_csr_t device_handle() const { return ptr_ ? ptr_->h : nullptr; }
};

}

#endif
