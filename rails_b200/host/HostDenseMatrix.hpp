// DenseMatrix concept for the CUDA backend: a small column-major host matrix.
//
// The projected k x k objects (VAV, VBV, T, H, the Ritz coefficient blocks)
// stay on the host by design: the solver hands their
// raw pointer + LDA() straight to sb03md / DSYEV
// (/root/reference/src/LyapunovSolver.hpp:357,458-459).  Semantics follow the
// reference's DenseMatrix implementations (StlWrapper used as DenseMatrix,
// /root/reference/src/StlWrapper.cpp; Epetra_SerialDenseMatrixWrapper.cpp:58-241):
// content-preserving capacity resize, shallow operator= unless the target is
// a view, deep copy(), transpose as a flag on a shallow handle.
#ifndef RAILS_B200_HOSTDENSEMATRIX_HPP
#define RAILS_B200_HOSTDENSEMATRIX_HPP

#include <memory>
#include <vector>

namespace RAILS
{

class HostDenseMatrix
{
    std::shared_ptr<std::vector<double> > buf_;
    int m_, n_;          // logical size
    int ld_, n_cap_;     // allocation: ld_ rows x n_cap_ columns
    bool is_view_;
    bool transpose_;

public:
    HostDenseMatrix();
    HostDenseMatrix(int m, int n);
    // deep copy, like every reference wrapper's copy constructor
    HostDenseMatrix(HostDenseMatrix const &other);
    HostDenseMatrix(HostDenseMatrix &&other) noexcept;

    HostDenseMatrix &operator=(HostDenseMatrix const &other);
    HostDenseMatrix &operator=(HostDenseMatrix &&other) noexcept;
    HostDenseMatrix &operator=(double value);

    HostDenseMatrix &operator*=(double value);
    HostDenseMatrix &operator/=(double value);

    HostDenseMatrix operator*(HostDenseMatrix const &other) const;

    double &operator()(int i, int j = 0);
    double const &operator()(int i, int j = 0) const;

    operator double *() const;

    void resize(int m, int n);

    double norm_inf() const;

    HostDenseMatrix view();
    HostDenseMatrix copy() const;
    HostDenseMatrix transpose() const;

    int M() const;
    int N() const;
    int LDA() const;

    // all eigenpairs of the symmetric matrix (DSYEV 'V','U'), optionally the
    // `num` largest in magnitude above `tol` (StlWrapper::eigs :433-479)
    int eigs(HostDenseMatrix &V, HostDenseMatrix &D, int num = -1, double tol = 1e-16) const;

    bool transposed() const { return transpose_; }
};

}

#endif
