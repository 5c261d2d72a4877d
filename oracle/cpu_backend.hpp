// TEST INFRASTRUCTURE (oracle/_ref build; never linked into the product).
//
// CsrHostOperator: the Operator concept of RAILS::Solver
// (/root/reference/src/LyapunovSolverDecl.hpp:9-51) over a host CSR matrix, applied
// with OpenMP over rows.  It is the ONE class the reference's CPU build lacks
// without Trilinos: its sparse operator is Epetra_OperatorWrapper
// (src/Epetra_OperatorWrapper.cpp:75-91 -> Epetra_CrsMatrix::Apply: y_i = sum_j a_ij x_j),
// and its Stl build stores A dense, which stops at m ~ 1e4.  With this
// operator the UNMODIFIED reference loop runs sparse problems of any size on the CPU
// with the reference's own StlWrapper as MultiVector and DenseMatrix:
//     Solver<CsrHostOperator, StlWrapper, StlWrapper>
// -- the timed CPU baseline of bench.py (`--impl reference`, kind "reference") and a
// second parity arm.  Member semantics follow Epetra_OperatorWrapper: shallow handle
// copy (:41-48), transpose() flips a flag on a copy (:50-55), M()/N() swap under it
// (:93-113).
#ifndef RAILS_ORACLE_CPU_BACKEND_HPP
#define RAILS_ORACLE_CPU_BACKEND_HPP

#include <memory>
#include <vector>

#include "src/StlWrapper.hpp"

namespace rails_oracle
{

class CsrHostOperator
{
    struct Data
    {
        int m, n;
        std::vector<int> rowptr, colidx;
        std::vector<double> vals;
    };
    std::shared_ptr<const Data> ptr_;
    bool transpose_;

public:
    CsrHostOperator();
    CsrHostOperator(int m, int n, const int *rowptr, const int *colidx, const double *vals);

    CsrHostOperator transpose() const;
    RAILS::StlWrapper operator*(RAILS::StlWrapper const &other) const;

    // 2-norm by power iteration on A'A (only reached when B is an Operator)
    double norm() const;

    int M() const;
    int N() const;
};

}

#endif
