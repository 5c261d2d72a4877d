// TEST INFRASTRUCTURE: see cpu_backend.hpp.
#include "cpu_backend.hpp"

#include <cmath>
#include <iostream>

namespace rails_oracle
{

CsrHostOperator::CsrHostOperator() : transpose_(false) {}

CsrHostOperator::CsrHostOperator(int m, int n, const int *rowptr, const int *colidx, const double *vals)
    : transpose_(false)
{
    auto d = std::make_shared<Data>();
    d->m = m;
    d->n = n;
    d->rowptr.assign(rowptr, rowptr + m + 1);
    d->colidx.assign(colidx, colidx + rowptr[m]);
    d->vals.assign(vals, vals + rowptr[m]);
    ptr_ = d;
}

CsrHostOperator CsrHostOperator::transpose() const
{
    CsrHostOperator out(*this);
    out.transpose_ = !out.transpose_;
    return out;
}

int CsrHostOperator::M() const { return !ptr_ ? -1 : (transpose_ ? ptr_->n : ptr_->m); }
int CsrHostOperator::N() const { return !ptr_ ? -1 : (transpose_ ? ptr_->m : ptr_->n); }

RAILS::StlWrapper CsrHostOperator::operator*(RAILS::StlWrapper const &other) const
{
    RAILS::StlWrapper out(M(), other.N());
    if (!ptr_ || other.M() != N())
    {
        std::cerr << "Incomplatible sizes: operator " << M() << "x" << N() << " times "
                  << other.M() << "x" << other.N() << std::endl;
        return out;
    }
    Data const &A = *ptr_;
    const double *X = other;           // column-major, leading dimension LDA()
    double *Y = out;
    const long long ldx = other.LDA(), ldy = out.LDA();
    const int cols = other.N();
    if (!transpose_)
    {
        // each row in CSR order with separately rounded multiply and add
        // (Epetra_CrsMatrix::Apply's loop; no FMA: the recipe compiles with -ffp-contract=off)
#pragma omp parallel for schedule(static)
        for (int i = 0; i < A.m; ++i)
            for (int c = 0; c < cols; ++c)
            {
                double s = 0.0;
                for (int p = A.rowptr[i]; p < A.rowptr[i + 1]; ++p)
                    s += A.vals[p] * X[A.colidx[p] + c * ldx];
                Y[i + c * ldy] = s;
            }
    }
    else
    {
#pragma omp parallel for schedule(static)
        for (int c = 0; c < cols; ++c)
        {
            double *y = Y + c * ldy;
            for (int j = 0; j < A.n; ++j) y[j] = 0.0;
            for (int i = 0; i < A.m; ++i)
            {
                const double x = X[i + c * ldx];
                for (int p = A.rowptr[i]; p < A.rowptr[i + 1]; ++p)
                    y[A.colidx[p]] += A.vals[p] * x;
            }
        }
    }
    return out;
}

double CsrHostOperator::norm() const
{
    if (!ptr_) return 0.0;
    CsrHostOperator plain(*this);
    plain.transpose_ = false;
    RAILS::StlWrapper x(ptr_->n, 1);
    x = 1.0;
    double lambda = 0.0, prev = -1.0;
    for (int it = 0; it < 1000; ++it)
    {
        const double nx = x.norm();
        if (nx == 0.0) return 0.0;
        x /= nx;
        RAILS::StlWrapper y = plain * x;
        lambda = y.dot(y)(0, 0);
        x = plain.transpose() * y;
        if (std::abs(lambda - prev) <= 1e-14 * std::abs(lambda)) break;
        prev = lambda;
    }
    return std::sqrt(lambda);
}

}
