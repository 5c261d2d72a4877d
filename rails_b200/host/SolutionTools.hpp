// Post-processing of a low-rank solution X = V T V' (SURVEY.md §8(f) N2).
//
// The reference's executable wraps X in an operator and asks Anasazi's block
// Krylov-Schur for its `num` dominant eigenpairs, then prints each eigenvalue
// next to its share of the trace (/root/reference/src/main.cpp:140-170,
// src/Epetra_OperatorWrapper.cpp:147-222, trace: src/SchurOperator.cpp:298-342).
// An iterative eigensolver is not needed for that: X has rank k, and with
// G = V'V = L L' (G = I for the orthonormal basis the solver returns)
//
//     X u = lambda u,  u = V c   <=>   (L' T L) w = lambda w,  c = L^-T w,
//
// so the eigenvalues are those of a k x k symmetric matrix (host, DSYEV) and
// the eigenvectors are ONE tall-skinny product V (L^-T W) on the device -- the
// same V*C kernel the solver uses.  trace(X) = trace(T G).
//
// Written against the MultiVector / DenseMatrix concepts of the solver, so it
// works for any backend; the CUDA wrappers run V'V and V*C as kernels.
#ifndef RAILS_B200_SOLUTIONTOOLS_HPP
#define RAILS_B200_SOLUTIONTOOLS_HPP

#include <algorithm>
#include <cmath>
#include <iostream>
#include <vector>

namespace RAILS
{
namespace b200
{

// Dominant eigenpairs of X = V T V'.  Returns at most `num` pairs (num < 1:
// all), ordered by decreasing magnitude, and stops at the first eigenvalue
// with |lambda| < tol like the reference's eigs()
// (src/Epetra_OperatorWrapper.cpp:200-212).  eigenvalues is r x 1,
// eigenvectors m x r with orthonormal columns.  Returns 0, or the LAPACK
// `info` of the k x k eigensolve / 1 when V'V is not positive definite.
template<class MultiVector, class DenseMatrix>
int solution_eigs(MultiVector const &V, DenseMatrix const &T, MultiVector &eigenvectors,
                  DenseMatrix &eigenvalues, int num = 10, double tol = 1e-6)
{
    const int k = V.N();
    if (k < 1 || T.M() != k || T.N() != k)
    {
        std::cerr << "solution_eigs: V has " << k << " columns, T is " << T.M() << "x" << T.N()
                  << std::endl;
        return 1;
    }
    if (num < 1 || num > k)
        num = k;

    // Gram matrix and its Cholesky factor (lower, column-major loops: k is small)
    DenseMatrix G = V.dot(V);
    double dev = 0.0;
    for (int j = 0; j < k; ++j)
        for (int i = 0; i < k; ++i)
            dev = std::max(dev, std::abs(G(i, j) - (i == j ? 1.0 : 0.0)));
    const bool orthonormal = dev < 1e-12;

    DenseMatrix L(k, k);
    L = 0.0;
    if (orthonormal)
        for (int i = 0; i < k; ++i)
            L(i, i) = 1.0;
    else
        for (int j = 0; j < k; ++j)
        {
            double d = G(j, j);
            for (int p = 0; p < j; ++p)
                d -= L(j, p) * L(j, p);
            if (!(d > 0.0))
            {
                std::cerr << "solution_eigs: V'V is not positive definite (column " << j << ")"
                          << std::endl;
                return 1;
            }
            L(j, j) = std::sqrt(d);
            for (int i = j + 1; i < k; ++i)
            {
                double s = 0.5 * (G(i, j) + G(j, i));
                for (int p = 0; p < j; ++p)
                    s -= L(i, p) * L(j, p);
                L(i, j) = s / L(j, j);
            }
        }

    // S = L' T L, symmetrised
    DenseMatrix S = L.transpose() * (T * L);
    for (int j = 0; j < k; ++j)
        for (int i = 0; i < j; ++i)
        {
            const double v = 0.5 * (S(i, j) + S(j, i));
            S(i, j) = v;
            S(j, i) = v;
        }

    DenseMatrix W(k, k), lambda(k, 1);
    int info = S.eigs(W, lambda);
    if (info)
        return info;

    std::vector<int> order(k);
    for (int i = 0; i < k; ++i)
        order[i] = i;
    std::stable_sort(order.begin(), order.end(), [&](int a, int b)
                     { return std::abs(lambda(a, 0)) > std::abs(lambda(b, 0)); });
    int kept = 0;
    while (kept < num && std::abs(lambda(order[kept], 0)) >= tol)
        ++kept;

    eigenvalues = DenseMatrix(std::max(kept, 1), 1);
    eigenvalues.resize(kept, 1);
    DenseMatrix C(k, std::max(kept, 1));
    C.resize(k, kept);
    for (int c = 0; c < kept; ++c)
    {
        eigenvalues(c, 0) = lambda(order[c], 0);
        // solve L' x = w by back substitution
        for (int i = k - 1; i >= 0; --i)
        {
            double s = W(i, order[c]);
            for (int p = i + 1; p < k; ++p)
                s -= L(p, i) * C(p, c);
            C(i, c) = s / L(i, i);
        }
    }
    if (kept > 0)
        eigenvectors = V * C;
    else
    {
        eigenvectors = MultiVector(V, 1);
        eigenvectors.resize(0);
    }
    return 0;
}

// trace(V T V') = trace(T V'V); `orthonormal` skips the Gram matrix (then it
// is the sum of T's diagonal, the first term of src/SchurOperator.cpp:311-313)
template<class MultiVector, class DenseMatrix>
double solution_trace(MultiVector const &V, DenseMatrix const &T, bool orthonormal = false)
{
    const int k = T.M();
    double trace = 0.0;
    if (orthonormal)
    {
        for (int i = 0; i < k; ++i)
            trace += T(i, i);
        return trace;
    }
    DenseMatrix G = V.dot(V);
    for (int j = 0; j < k; ++j)
        for (int i = 0; i < k; ++i)
            trace += T(i, j) * G(j, i);
    return trace;
}

}
}

#endif
