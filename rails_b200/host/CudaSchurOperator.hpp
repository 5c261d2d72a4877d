// Schur-complement operator for a singular (diagonal) mass matrix on the B200
// backend: the counterpart of RAILS::SchurOperator
// (/root/reference/src/SchurOperator.hpp:23-88, .cpp:51-342) and of MATLAB's
// RAILSschur (matlab/RAILSschur.m:1-92).
//
// With idx1 = rows where M is zero and idx2 the rest, the Lyapunov equation
// A X M' + M X A' + B B' = 0 reduces to the non-singular part
//     S X22 MS' + MS X22 S' + BS BS' = 0,    S = A22 - A21 A11^-1 A12,
// MS = M(idx2, idx2), BS = B(idx2, :) (or its restriction when B is not zero in
// the singular part), and the full solution is X = P X22 P' with the
// prolongation P = [-A11^-1 A12; I] (RAILSschur.m:74-76).
//
// Set-up (Compute, once, host): split A into its four blocks, factorise A11
// with the sparse LU of SparseLU.hpp, upload blocks and factors.
// Apply (every iteration, device): two rectangular SpMMs, one LU solve (four
// permutation/triangular sweeps), one SpMM, one axpy -- all through the C ABI of
// include/rails_cuda.h.  Single GPU: the reference's KLU solve is serial as well
// and its distributed Apply is marked "TODO: fix these maps to work in
// parallel" (src/SchurOperator.cpp:229).
#ifndef RAILS_B200_CUDASCHUROPERATOR_HPP
#define RAILS_B200_CUDASCHUROPERATOR_HPP

#include "CudaMultiVectorWrapper.hpp"
#include "CudaOperatorWrapper.hpp"
#include "HostDenseMatrix.hpp"

#include <memory>
#include <vector>

namespace RAILS
{

class CudaSchurOperator
{
public:
    struct State;

    CudaSchurOperator() {}
    // A: n x n host CSR; mdiag: the n diagonal entries of M (the reference
    // extracts them the same way, src/SchurOperator.cpp:56-60); rows with
    // |M_ii| < zero_tolerance form the singular part (1e-12 in
    // matlab/RAILSschur.m:23-24, 1e-15 in src/SchurOperator.cpp:75).
    // `scale` multiplies A first: the C++ reference passes -1 because its solver
    // has no mass matrix and its M22 is -I (src/SchurOperator.cpp:134-135); with
    // the generalised loop ("Use mass matrix") leave it at 1.
    CudaSchurOperator(int n, const int *rowptr, const int *colidx, const double *vals,
                      const double *mdiag, double zero_tolerance = 1e-12, double scale = 1.0);

    // splits, factorises and uploads; 0 on success (SchurOperator::Compute, :51-188)
    int Compute();

    int n() const;       // rows of the original problem
    int n1() const;      // singular part
    int n2() const;      // size of the Schur complement
    std::vector<int> const &idx1() const;
    std::vector<int> const &idx2() const;

    // Y = S X  (SchurOperator::Apply without a solution, :190-219;
    // AS / ASt of RAILSschur.m:35-36)
    CudaMultiVectorWrapper apply(CudaMultiVectorWrapper const &X, bool transpose = false) const;
    // Y = S^-1 X = (A^-1 [0; X])(idx2)  (Sinv, RAILSschur.m:60-66).  The first
    // call factorises the whole of A.
    CudaMultiVectorWrapper apply_inverse(CudaMultiVectorWrapper const &X) const;

    // the three operators the solver takes
    CudaOperatorWrapper schur_operator() const;      // S
    CudaOperatorWrapper inverse_operator() const;    // Sinv
    CudaOperatorWrapper mass_operator() const;       // MS = M(idx2, idx2)
    bool mass_is_identity() const;

    // BS: B(idx2, :), or B(idx2, :) - A21 A11^-1 B(idx1, :) when B is not zero in
    // the singular part (RAILSschur.m:44-49,70-72)
    CudaMultiVectorWrapper restrict(CudaMultiVectorWrapper const &B) const;
    // Vtrans: n2 rows -> n rows, [-A11^-1 A12 V; V] in the original ordering
    // (RAILSschur.m:74-76).  X = Vfull T Vfull' is the solution of the original
    // problem, which is what SchurOperator::Apply with SetSolution multiplies by
    // (:222-293): its blocks X22 = V T V', X12 = -A11^-1 A12 X22, X21 = X12',
    // X11 = A11^-1 A12 X22 A12' A11^-T are exactly P V T V' P'.
    CudaMultiVectorWrapper prolongate(CudaMultiVectorWrapper const &V) const;
    // rows idx2 of a full-length block: the reduced basis back from a prolongated
    // one (prolongate leaves V itself in those rows), X2 of RAILSschur.m:58
    CudaMultiVectorWrapper select_regular(CudaMultiVectorWrapper const &full) const;

    // trace of the full solution X = P V T V' P' (SchurOperator::Trace, :295-342):
    // trace(T) + trace(T V' A12' A11^-T A11^-1 A12 V) for an orthonormal V
    double Trace(CudaMultiVectorWrapper const &V, HostDenseMatrix const &T) const;

    // operator applies so far, counted per column (SchurOperator::GetMVPs, :396)
    int GetMVPs() const;

    // set-up facts for the driver's report
    long long factor_nnz() const;
    int factor_levels(int which) const;
    int launches_per_solve() const;
    int off_diagonal_pivots() const;
    double factor_seconds() const;

private:
    std::shared_ptr<State> state_;
};

}

#endif
