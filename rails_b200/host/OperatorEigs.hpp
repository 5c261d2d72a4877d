// Largest-magnitude eigenpairs of a symmetric operator by Lanczos with full
// re-orthogonalisation, written against the wrapper concepts only (Operator *
// MultiVector, dot, view, operator-=, norm): the stand-in for the Anasazi block
// Krylov-Schur solve behind the reference's Operator::eigs
// (/root/reference/src/Epetra_OperatorWrapper.cpp:147-222).  Not on the solver's
// path; the typed Operator suite exercises it (test/GenericOperatorWrapper_test.cpp:116-185).
#ifndef RAILS_B200_OPERATOREIGS_HPP
#define RAILS_B200_OPERATOREIGS_HPP

#include <algorithm>
#include <cmath>
#include <iostream>
#include <vector>

namespace RAILS
{

// Start vector: all ones, as the reference (:172-174).  The Krylov space grows until
// the `num` Ritz pairs of largest magnitude have residuals |beta y_last| below the
// tolerance, or the space is exhausted.  Eigenvalues whose magnitude is below `tol`
// are cut off after the solve, as the reference does (:203-214).
template<class Operator, class MultiVector, class DenseMatrix>
int symmetric_eigs(Operator const &A, MultiVector &V, DenseMatrix &D, int num, double tol)
{
    const int n = A.N();
    if (A.M() != n || n < 1 || num < 1)
    {
        std::cerr << "eigs needs a square operator and num >= 1" << std::endl;
        return 1;
    }
    num = std::min(num, n);
    const int max_dim = std::min(n, std::max(4 * num + 40, 60));
    MultiVector Q(n, max_dim + 1);
    Q.resize(1);
    Q = 1.0;
    Q /= Q.norm();
    std::vector<double> alpha, beta;
    DenseMatrix Y, theta;
    int dim = 0;
    while (dim < max_dim)
    {
        MultiVector w = A * Q.view(dim);
        // two Gram-Schmidt passes against the whole basis (the projection on the
        // current vector is the diagonal entry of the tridiagonal matrix)
        double a = 0.0;
        for (int pass = 0; pass < 2; ++pass)
        {
            DenseMatrix c = Q.dot(w);
            a += c(dim, 0);
            w -= Q * c;
        }
        alpha.push_back(a);
        const double b = w.norm();
        dim++;
        double scale = std::abs(a);
        for (double x : beta) scale = std::max(scale, x);
        const bool broke_down = !(b > 1e-13 * std::max(scale, 1e-300));
        const bool exhausted = broke_down || dim == max_dim;

        // Ritz pairs of the dim x dim tridiagonal matrix
        DenseMatrix H(dim, dim);
        H = 0.0;
        for (int i = 0; i < dim; ++i)
        {
            H(i, i) = alpha[i];
            if (i + 1 < dim) { H(i + 1, i) = beta[i]; H(i, i + 1) = beta[i]; }
        }
        H.eigs(Y, theta);
        std::vector<int> order(dim);
        for (int i = 0; i < dim; ++i) order[i] = i;
        std::sort(order.begin(), order.end(),
                  [&](int x, int y) { return std::abs(theta(x, 0)) > std::abs(theta(y, 0)); });
        // `tol` is first of all the cut-off below (the solver passes its restart
        // tolerance); the Ritz residuals are held to the tighter of tol and 1e-10 of
        // the largest Ritz value, so that a generous cut-off does not return
        // half-converged values
        const double residual_tol = std::min(tol, 1e-10 * std::abs(theta(order[0], 0)));
        bool converged = dim >= num;
        for (int i = 0; i < std::min(num, dim) && converged; ++i)
            converged = std::abs(b * Y(dim - 1, order[i])) < residual_tol;
        if (converged || exhausted)
        {
            int kept = 0;
            const int want = std::min(num, dim);
            DenseMatrix C(dim, want);
            D.resize(want, 1);
            for (int i = 0; i < want; ++i)
            {
                if (std::abs(theta(order[i], 0)) < tol)
                    break;
                D(i, 0) = theta(order[i], 0);
                for (int r = 0; r < dim; ++r)
                    C(r, i) = Y(r, order[i]);
                kept++;
            }
            D.resize(kept, 1);
            C.resize(dim, kept);
            Q.resize(dim);
            V = Q * C;
            // an exhausted Krylov space holds exact eigenpairs of what it spans
            const bool ok = converged || broke_down || dim == n;
            if (!ok)
                std::cerr << "Eigensolver did not converge" << std::endl;
            return ok ? 0 : 1;
        }
        beta.push_back(b);
        Q.resize(dim + 1);
        Q.view(dim) = w;
        Q.view(dim) /= b;
    }
    return 1;
}

}

#endif
