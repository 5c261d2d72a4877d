#include "CudaSchurOperator.hpp"

#include "SparseLU.hpp"
#include "rails_cuda.h"

#include <chrono>
#include <cmath>
#include <iostream>

namespace RAILS
{

namespace
{

bool ok(int rc, const char *what)
{
    if (rc)
        std::cerr << "CudaSchurOperator::" << what << " failed (" << rc << "): "
                  << rails_cuda_last_error() << std::endl;
    return rc == 0;
}

struct HostCsr
{
    int rows = 0, cols = 0;
    std::vector<int> rowptr, colidx;
    std::vector<double> vals;
};

struct LuHandle
{
    rails_lu_t h = nullptr;
    ~LuHandle() { rails_lu_free(h); }
};

struct IdxHandle
{
    rails_idx_t h = nullptr;
    ~IdxHandle() { rails_idx_free(h); }
};

int upload_factors(SparseLU const &lu, LuHandle &out)
{
    return rails_lu_create(lu.n, lu.Lp.data(), lu.Li.data(), lu.Lx.data(), lu.Up.data(),
                           lu.Ui.data(), lu.Ux.data(), lu.pinv.data(), lu.q.data(), &out.h);
}

}

struct CudaSchurOperator::State
{
    int n = 0;
    double zero_tolerance = 1e-12;
    HostCsr A;                      // scaled copy of the input
    std::vector<double> mdiag;
    std::vector<int> idx1, idx2;
    CudaOperatorWrapper A12, A21, A22, MS;
    bool mass_is_identity = true;
    LuHandle lu11, lu_full;
    IdxHandle dev_idx1, dev_idx2;
    SparseLU host11;                // kept for the report (fill, pivots)
    double factor_seconds = 0.0;
    int mvps = 0;
    bool computed = false;

    int factor_full();
};

namespace
{

// A(rows_sel, cols_sel) with renumbered columns; col_new[j] < 0: not selected
HostCsr extract(HostCsr const &A, std::vector<int> const &rows_sel, std::vector<int> const &col_new,
                int ncols)
{
    HostCsr out;
    out.rows = (int)rows_sel.size();
    out.cols = ncols;
    out.rowptr.assign(1, 0);
    for (size_t r = 0; r < rows_sel.size(); ++r)
    {
        const int i = rows_sel[r];
        for (int p = A.rowptr[i]; p < A.rowptr[i + 1]; ++p)
        {
            const int j = col_new[A.colidx[p]];
            if (j >= 0)
            {
                out.colidx.push_back(j);
                out.vals.push_back(A.vals[p]);
            }
        }
        out.rowptr.push_back((int)out.colidx.size());
    }
    return out;
}

CudaOperatorWrapper to_device(HostCsr const &A)
{
    // empty blocks still need valid pointers
    static const int zero_i = 0;
    static const double zero_d = 0.0;
    return CudaOperatorWrapper(A.rows, A.rows, A.cols, 0, A.rowptr.data(),
                               A.colidx.empty() ? &zero_i : A.colidx.data(),
                               A.vals.empty() ? &zero_d : A.vals.data());
}

struct SchurAction : CudaOperatorWrapper::Action
{
    CudaSchurOperator op;
    bool inverse;
    SchurAction(CudaSchurOperator const &o, bool inv) : op(o), inverse(inv) {}
    CudaMultiVectorWrapper apply(CudaMultiVectorWrapper const &x, bool transpose) const
    {
        if (inverse)
        {
            if (transpose)
                std::cerr << "CudaSchurOperator: the transposed inverse is not implemented" << std::endl;
            return op.apply_inverse(x);
        }
        return op.apply(x, transpose);
    }
    int rows() const { return op.n2(); }
    int cols() const { return op.n2(); }
};

}

CudaSchurOperator::CudaSchurOperator(int n, const int *rowptr, const int *colidx, const double *vals,
                                     const double *mdiag, double zero_tolerance, double scale)
    : state_(std::make_shared<State>())
{
    State &s = *state_;
    s.n = n;
    s.zero_tolerance = zero_tolerance;
    s.A.rows = s.A.cols = n;
    s.A.rowptr.assign(rowptr, rowptr + n + 1);
    s.A.colidx.assign(colidx, colidx + rowptr[n]);
    s.A.vals.assign(vals, vals + rowptr[n]);
    if (scale != 1.0)
        for (size_t p = 0; p < s.A.vals.size(); ++p)
            s.A.vals[p] *= scale;
    s.mdiag.assign(mdiag, mdiag + n);
}

int CudaSchurOperator::Compute()
{
    if (!state_)
        return 1;
    State &s = *state_;
    if (rails_cuda_comm_size() > 1)
    {
        std::cerr << "CudaSchurOperator: the Schur complement runs on one GPU" << std::endl;
        return 1;
    }

    // split on the diagonal of M (src/SchurOperator.cpp:71-79, RAILSschur.m:23-24)
    std::vector<int> new1(s.n, -1), new2(s.n, -1);
    s.idx1.clear();
    s.idx2.clear();
    for (int i = 0; i < s.n; ++i)
    {
        if (std::abs(s.mdiag[i]) < s.zero_tolerance)
        {
            new1[i] = (int)s.idx1.size();
            s.idx1.push_back(i);
        }
        else
        {
            new2[i] = (int)s.idx2.size();
            s.idx2.push_back(i);
        }
    }
    const int n1 = (int)s.idx1.size(), n2 = (int)s.idx2.size();
    if (n2 == 0)
    {
        std::cerr << "CudaSchurOperator: M is zero, nothing to solve" << std::endl;
        return 1;
    }

    HostCsr A11 = extract(s.A, s.idx1, new1, n1);
    HostCsr A12 = extract(s.A, s.idx1, new2, n2);
    HostCsr A21 = extract(s.A, s.idx2, new1, n1);
    HostCsr A22 = extract(s.A, s.idx2, new2, n2);

    // factorise A11 on the host (Amesos_Klu Symbolic/NumericFactorization, :185-186)
    auto t0 = std::chrono::steady_clock::now();
    if (n1 > 0)
    {
        int info = s.host11.factor(n1, A11.rowptr.data(), A11.colidx.data(), A11.vals.data());
        if (info)
        {
            std::cerr << "CudaSchurOperator: A11 is singular (no pivot at step " << info - 1 << ")"
                      << std::endl;
            return 2;
        }
    }
    s.factor_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();

    if (n1 > 0)
    {
        if (!ok(upload_factors(s.host11, s.lu11), "Compute (factors of A11)"))
            return 3;
        s.A12 = to_device(A12);
        s.A21 = to_device(A21);
        if (!s.A12.valid() || !s.A21.valid())
            return 3;
    }
    s.A22 = to_device(A22);
    if (!s.A22.valid())
        return 3;

    // MS = M(idx2, idx2) as a diagonal CSR matrix
    HostCsr MS;
    MS.rows = MS.cols = n2;
    MS.rowptr.resize(n2 + 1);
    MS.colidx.resize(n2);
    MS.vals.resize(n2);
    s.mass_is_identity = true;
    for (int i = 0; i < n2; ++i)
    {
        MS.rowptr[i] = i;
        MS.colidx[i] = i;
        MS.vals[i] = s.mdiag[s.idx2[i]];
        if (MS.vals[i] != 1.0)
            s.mass_is_identity = false;
    }
    MS.rowptr[n2] = n2;
    s.MS = to_device(MS);
    if (!s.MS.valid())
        return 3;

    if (!ok(rails_idx_create(n1, s.idx1.data(), &s.dev_idx1.h), "Compute (idx1)") ||
        !ok(rails_idx_create(n2, s.idx2.data(), &s.dev_idx2.h), "Compute (idx2)"))
        return 3;
    s.computed = true;
    return 0;
}

int CudaSchurOperator::State::factor_full()
{
    if (lu_full.h)
        return 0;
    // [L, U, P, Q] = lu(A) of the `factorize` option (RAILSschur.m:62-65)
    SparseLU lu;
    auto t0 = std::chrono::steady_clock::now();
    int info = lu.factor(n, A.rowptr.data(), A.colidx.data(), A.vals.data());
    factor_seconds += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (info)
    {
        std::cerr << "CudaSchurOperator: A is singular (no pivot at step " << info - 1 << ")" << std::endl;
        return 2;
    }
    return ok(upload_factors(lu, lu_full), "inverse (factors of A)") ? 0 : 3;
}

int CudaSchurOperator::n() const { return state_ ? state_->n : 0; }
int CudaSchurOperator::n1() const { return state_ ? (int)state_->idx1.size() : 0; }
int CudaSchurOperator::n2() const { return state_ ? (int)state_->idx2.size() : 0; }
std::vector<int> const &CudaSchurOperator::idx1() const { return state_->idx1; }
std::vector<int> const &CudaSchurOperator::idx2() const { return state_->idx2; }
bool CudaSchurOperator::mass_is_identity() const { return state_ && state_->mass_is_identity; }
int CudaSchurOperator::GetMVPs() const { return state_ ? state_->mvps : 0; }
long long CudaSchurOperator::factor_nnz() const { return state_ ? rails_lu_nnz(state_->lu11.h) : 0; }
int CudaSchurOperator::factor_levels(int which) const { return state_ ? rails_lu_levels(state_->lu11.h, which) : 0; }
int CudaSchurOperator::launches_per_solve() const { return state_ ? rails_lu_launches_per_solve(state_->lu11.h, 0) : 0; }
int CudaSchurOperator::off_diagonal_pivots() const { return state_ ? state_->host11.off_diagonal_pivots : 0; }
double CudaSchurOperator::factor_seconds() const { return state_ ? state_->factor_seconds : 0.0; }

CudaMultiVectorWrapper CudaSchurOperator::apply(CudaMultiVectorWrapper const &X, bool transpose) const
{
    State &s = *state_;
    const int c = X.N();
    if (!s.computed || X.M() != n2())
    {
        std::cerr << "CudaSchurOperator::apply: operator is " << n2() << "x" << n2()
                  << (s.computed ? "" : " (Compute() has not run)") << ", X has " << X.M()
                  << " rows" << std::endl;
        return CudaMultiVectorWrapper(n2(), c);
    }
    s.mvps += c;
    // Y = A22 X - A21 (A11 \ (A12 X))                      (:195-217)
    // Y = A22' X - A12' (A11' \ (A21' X))   transposed     (RAILSschur.m:36)
    CudaMultiVectorWrapper Y = (transpose ? s.A22.transpose() : s.A22) * X;
    if (n1() == 0 || c == 0)
        return Y;
    CudaMultiVectorWrapper t = (transpose ? s.A21.transpose() : s.A12) * X;
    ok(rails_lu_solve(s.lu11.h, transpose ? 1 : 0, t.device_handle(), t.device_col0(), c,
                      t.device_handle(), t.device_col0()), "apply (A11 solve)");
    Y -= (transpose ? s.A12.transpose() : s.A21) * t;
    return Y;
}

CudaMultiVectorWrapper CudaSchurOperator::apply_inverse(CudaMultiVectorWrapper const &X) const
{
    State &s = *state_;
    const int c = X.N();
    CudaMultiVectorWrapper Y(n2(), c);
    if (!s.computed || X.M() != n2() || s.factor_full())
    {
        std::cerr << "CudaSchurOperator::apply_inverse: not available" << std::endl;
        return Y;
    }
    if (c == 0)
        return Y;
    // (A \ Xreorder([0; x]))(idx2)                          (RAILSschur.m:60)
    CudaMultiVectorWrapper full(s.n, c);
    full = 0.0;
    ok(rails_mv_scatter_rows(full.device_handle(), full.device_col0(), X.device_handle(),
                             X.device_col0(), c, s.dev_idx2.h, 1.0), "apply_inverse (scatter)");
    ok(rails_lu_solve(s.lu_full.h, 0, full.device_handle(), full.device_col0(), c,
                      full.device_handle(), full.device_col0()), "apply_inverse (solve)");
    ok(rails_mv_gather_rows(Y.device_handle(), Y.device_col0(), full.device_handle(),
                            full.device_col0(), c, s.dev_idx2.h, 1.0), "apply_inverse (gather)");
    return Y;
}

CudaOperatorWrapper CudaSchurOperator::schur_operator() const
{
    return CudaOperatorWrapper(std::shared_ptr<const CudaOperatorWrapper::Action>(
        new SchurAction(*this, false)));
}

CudaOperatorWrapper CudaSchurOperator::inverse_operator() const
{
    return CudaOperatorWrapper(std::shared_ptr<const CudaOperatorWrapper::Action>(
        new SchurAction(*this, true)));
}

CudaOperatorWrapper CudaSchurOperator::mass_operator() const { return state_->MS; }

CudaMultiVectorWrapper CudaSchurOperator::restrict(CudaMultiVectorWrapper const &B) const
{
    State &s = *state_;
    const int c = B.N();
    CudaMultiVectorWrapper BS(n2(), c);
    if (!s.computed || B.M() != s.n)
    {
        std::cerr << "CudaSchurOperator::restrict: B has " << B.M() << " rows, expected " << s.n << std::endl;
        return BS;
    }
    if (c == 0)
        return BS;
    ok(rails_mv_gather_rows(BS.device_handle(), BS.device_col0(), B.device_handle(), B.device_col0(),
                            c, s.dev_idx2.h, 1.0), "restrict (gather idx2)");
    if (n1() == 0)
        return BS;
    CudaMultiVectorWrapper B1(n1(), c);
    ok(rails_mv_gather_rows(B1.device_handle(), B1.device_col0(), B.device_handle(), B.device_col0(),
                            c, s.dev_idx1.h, 1.0), "restrict (gather idx1)");
    if (B1.norm_inf() > std::sqrt(2.220446049250313e-16))          // RAILSschur.m:46-49
    {
        std::cerr << "Warning: B is not zero in the singular part" << std::endl;
        ok(rails_lu_solve(s.lu11.h, 0, B1.device_handle(), B1.device_col0(), c, B1.device_handle(),
                          B1.device_col0()), "restrict (A11 solve)");
        BS -= s.A21 * B1;
    }
    return BS;
}

CudaMultiVectorWrapper CudaSchurOperator::prolongate(CudaMultiVectorWrapper const &V) const
{
    State &s = *state_;
    const int c = V.N();
    CudaMultiVectorWrapper out(s.n, c);
    if (!s.computed || V.M() != n2())
    {
        std::cerr << "CudaSchurOperator::prolongate: V has " << V.M() << " rows, expected " << n2() << std::endl;
        return out;
    }
    if (c == 0)
        return out;
    ok(rails_mv_scatter_rows(out.device_handle(), out.device_col0(), V.device_handle(), V.device_col0(),
                             c, s.dev_idx2.h, 1.0), "prolongate (scatter idx2)");
    if (n1() == 0)
        return out;
    CudaMultiVectorWrapper t = s.A12 * V;
    ok(rails_lu_solve(s.lu11.h, 0, t.device_handle(), t.device_col0(), c, t.device_handle(),
                      t.device_col0()), "prolongate (A11 solve)");
    ok(rails_mv_scatter_rows(out.device_handle(), out.device_col0(), t.device_handle(), t.device_col0(),
                             c, s.dev_idx1.h, -1.0), "prolongate (scatter idx1)");
    return out;
}

CudaMultiVectorWrapper CudaSchurOperator::select_regular(CudaMultiVectorWrapper const &full) const
{
    State &s = *state_;
    const int c = full.N();
    CudaMultiVectorWrapper out(n2(), c);
    if (!s.computed || full.M() != s.n)
    {
        std::cerr << "CudaSchurOperator::select_regular: block has " << full.M() << " rows, expected "
                  << s.n << std::endl;
        return out;
    }
    if (c > 0)
        ok(rails_mv_gather_rows(out.device_handle(), out.device_col0(), full.device_handle(),
                                full.device_col0(), c, s.dev_idx2.h, 1.0), "select_regular");
    return out;
}

double CudaSchurOperator::Trace(CudaMultiVectorWrapper const &V, HostDenseMatrix const &T) const
{
    // trace(P V T V' P') = trace(T (V'V + W'W)),  W = A11^-1 A12 V; the reference
    // assumes V'V = I (trace(T) + trace(T V' A12' A11^-T A11^-1 A12 V), :308-341)
    State &s = *state_;
    const int k = V.N();
    if (!s.computed || k < 1 || T.M() < k || T.N() < k)
        return 0.0;
    HostDenseMatrix G = V.dot(V);
    if (n1() > 0)
    {
        CudaMultiVectorWrapper W = s.A12 * V;
        ok(rails_lu_solve(s.lu11.h, 0, W.device_handle(), W.device_col0(), k, W.device_handle(),
                          W.device_col0()), "Trace (A11 solve)");
        HostDenseMatrix G1 = W.dot(W);
        for (int j = 0; j < k; ++j)
            for (int i = 0; i < k; ++i)
                G(i, j) += G1(i, j);
    }
    double trace = 0.0;
    for (int i = 0; i < k; ++i)
        for (int j = 0; j < k; ++j)
            trace += T(i, j) * G(j, i);
    return trace;
}

}
