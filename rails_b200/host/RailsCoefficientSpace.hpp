// "Residual method" = 2: the RAILS iteration carried out in the coordinates of
// ONE orthonormal basis Q of span[B, V, AV(, MV)].
//
// The reference loop (src/LyapunovSolver.hpp:100-346) keeps V, AV (and BV) as
// separate tall-skinny blocks and, per outer iteration, streams them for the
// projection (:138-207), 2L times for the Lanczos estimate (:367-447) and four
// times for CGS2 (:339-340).  With the factorisation of "Residual method" = 1,
//
//     [B, V, AV, MV] = Q [R_B, R_V, R_A, R_M],      Q'Q = I,  r = rank columns,
//
// every one of those products is a small host product of the R blocks:
//
//     V'AV = R_V' R_A      V'BB'V = (R_B' R_V)' (R_B' R_V)      V'MV = R_V' R_M
//     residual  = Q (R_A T R_M' + R_M T R_A' + R_B R_B') Q'     (R_M = R_V for M = I)
//     new direction u = Q y,   V'u = R_V' y   =>   CGS2 of u against V happens on y
//     restart   V <- V X  is  R_V <- R_V X  (likewise R_A, R_M)
//
// so the only m-sized work of an iteration is: materialise the e new basis
// vectors W = Q Y (one pass over Q), apply the operator(s) to them, and
// orthogonalise A W (M W) against Q (tsqr_append: four passes over Q).  That is
// ~5 passes over r ~ k + n + e columns, against ~(2L + 7) passes over 2k + n
// columns in the reference's sequence.  In exact arithmetic the iterates are
// those of "Residual method" = 1 with "Orthogonalize candidates"
// (matlab/RAILSsolver.m:388-409: exact residual eigenpairs, candidates
// orthogonalised against V before they are appended); V itself is formed once,
// at the end, as Q R_V.
//
// Not available with a null space or M-orthogonalisation, or when B is an
// operator: solve() then runs "Residual method" = 1.
#ifndef RAILS_B200_RAILSCOEFFICIENTSPACE_HPP
#define RAILS_B200_RAILSCOEFFICIENTSPACE_HPP

extern "C" {
void dgeqp3_(int *m, int *n, double *a, int *lda, int *jpvt, double *tau, double *work,
             int *lwork, int *info);
void dorgqr_(int *m, int *n, int *k, double *a, int *lda, double *tau, double *work,
             int *lwork, int *info);
}

namespace RAILS
{
namespace b200
{

// After a restart the basis spans more than the compressed factor needs.
// Rank-revealing QR (column pivoting) of the r x c coefficient block
// [R_B, R_V, R_A, R_M] = Z S gives the r' <= c directions still in use:
// Q <- Q Z (one pass over Q), R blocks <- Z' R blocks.
namespace detail
{
// dst <- src inside dst's own allocation (resize keeps the capacity)
template<class DenseMatrix>
void assign_in_place(DenseMatrix &dst, DenseMatrix const &src)
{
    dst.resize(src.M(), src.N());
    for (int j = 0; j < src.N(); j++)
        for (int i = 0; i < src.M(); i++)
            dst(i, j) = src(i, j);
}
}

template<class Matrix, class MultiVector, class DenseMatrix>
void Solver<Matrix, MultiVector, DenseMatrix>::tsqr_shrink(TsqrState &st)
{
    DenseMatrix *blocks[4] = {&st.RB, &st.RV, &st.RA, &st.RM};
    int r = st.Q.N(), c = 0;
    for (int b = 0; b < 4; b++)
        c += blocks[b]->N();
    if (r <= c || c < 1)
        return;

    std::vector<double> all((size_t)r * c);
    int at = 0;
    for (int b = 0; b < 4; b++)
        for (int j = 0; j < blocks[b]->N(); j++, at++)
            for (int i = 0; i < r; i++)
                all[(size_t)i + (size_t)at * r] = (*blocks[b])(i, j);

    std::vector<int> pivot(c, 0);
    std::vector<double> tau(std::min(r, c));
    int info = 0, lwork = -1;
    double query = 0.0;
    dgeqp3_(&r, &c, all.data(), &r, pivot.data(), tau.data(), &query, &lwork, &info);
    lwork = std::max((int)query, 3 * c + 1);
    std::vector<double> work(lwork);
    dgeqp3_(&r, &c, all.data(), &r, pivot.data(), tau.data(), work.data(), &lwork, &info);
    if (info)
        return;
    const double top = std::abs(all[0]);
    int rank = 0;
    while (rank < std::min(r, c) && std::abs(all[(size_t)rank + (size_t)rank * r]) > 1e-13 * top)
        rank++;
    if (rank < 1 || rank >= r)
        return;
    lwork = -1;
    dorgqr_(&r, &rank, &rank, all.data(), &r, tau.data(), &query, &lwork, &info);
    lwork = std::max((int)query, rank);
    work.resize(lwork);
    dorgqr_(&r, &rank, &rank, all.data(), &r, tau.data(), work.data(), &lwork, &info);
    if (info)
        return;

    DenseMatrix Z(r, rank);
    for (int j = 0; j < rank; j++)
        for (int i = 0; i < r; i++)
            Z(i, j) = all[(size_t)i + (size_t)j * r];

    MultiVector shrunk = st.Q * Z;
    st.Q = shrunk;
    st.capacity = std::min(st.Q.M(), std::max(st.capacity, rank + 8));
    st.Q.resize(st.capacity);
    st.Q.resize(rank);
    for (int b = 0; b < 4; b++)
    {
        if (blocks[b]->N() < 1)
        {
            blocks[b]->resize(rank, 0);
            continue;
        }
        DenseMatrix reduced = Z.transpose() * (*blocks[b]);
        detail::assign_in_place(*blocks[b], reduced);
    }
}

template<class Matrix, class MultiVector, class DenseMatrix>
int Solver<Matrix, MultiVector, DenseMatrix>::solve_coefficient_space(MultiVector &V, DenseMatrix &T)
{
    const auto t_begin = std::chrono::steady_clock::now();
    history_.clear();
    sections_.clear();

    const int n = V.M();
    MultiVector const &B = B_.multivector();
    const int nb = B.N();
    const bool mass = use_mass_;

    if (!restart_from_solution_)
    {
        V.resize(1);
        V.random();                       // same draw as the reference (:108-115)
    }

    TsqrState st;
    st.capacity = std::min(n, std::max(100, 2 * V.N() + nb + 2 * expand_size_ + 2));
    st.Q = MultiVector(V, st.capacity);
    st.Q.resize(0);
    const int kcap = std::max(V.N(), std::min(restart_size_ > 0 ? restart_size_ : 100, n)) + expand_size_;
    st.RV = DenseMatrix(st.capacity, kcap); st.RV.resize(0, 0);
    st.RA = DenseMatrix(st.capacity, kcap); st.RA.resize(0, 0);
    st.RM = DenseMatrix(st.capacity, kcap); st.RM.resize(0, 0);
    st.RB = DenseMatrix(st.capacity, std::max(nb, 1)); st.RB.resize(0, 0);
    st.valid = true;

    typedef std::vector<MultiVector const *> Blocks;
    typedef std::vector<DenseMatrix *> Factors;
    {
        Section s(sections_, "orthogonalize");
        MultiVector const &Ball = B.view(0, nb - 1);
        tsqr_append(st, Blocks(1, &Ball), Factors(1, &st.RB));
        // the start basis: its coefficient columns are orthonormalised on the
        // host (two Gram-Schmidt passes), V = Q R_V from here on
        MultiVector const &Vall = static_cast<MultiVector const &>(V).view(0, V.N() - 1);
        tsqr_append(st, Blocks(1, &Vall), Factors(1, &st.RV));
    }
    // CGS2 of coefficient column `col` of X against its columns [0, upto) and,
    // when other is given, against all columns of *other; returns the remaining norm
    auto orthogonalize_column = [](DenseMatrix &X, int col, int upto, DenseMatrix const *other) {
        const int rows = X.M();
        for (int pass = 0; pass < 2; pass++)
        {
            for (int src = 0; src < 2; src++)
            {
                DenseMatrix const *Y = src == 0 ? other : &X;
                const int count = src == 0 ? (other ? other->N() : 0) : upto;
                if (!Y || count < 1)
                    continue;
                std::vector<double> c(count, 0.0);
                for (int j = 0; j < count; j++)
                    for (int i = 0; i < rows; i++)
                        c[j] += (*Y)(i, j) * X(i, col);
                for (int j = 0; j < count; j++)
                    for (int i = 0; i < rows; i++)
                        X(i, col) -= c[j] * (*Y)(i, j);
            }
        }
        double nrm = 0.0;
        for (int i = 0; i < rows; i++)
            nrm += X(i, col) * X(i, col);
        return std::sqrt(nrm);
    };
    {
        int kept = 0;
        const int given = st.RV.N(), rows = st.RV.M();
        for (int j = 0; j < given; j++)
        {
            if (kept != j)
                for (int i = 0; i < rows; i++)
                    st.RV(i, kept) = st.RV(i, j);
            const double nrm = orthogonalize_column(st.RV, kept, kept, nullptr);
            if (!(nrm > 1e-8))
                continue;
            for (int i = 0; i < rows; i++)
                st.RV(i, kept) /= nrm;
            kept++;
        }
        st.RV.resize(rows, kept);
    }

    // W = Q * (columns [from, R_V.N()) of R_V), A W and M W appended to the basis
    auto apply_operators = [&](int from) {
        const int k = st.RV.N(), r = st.Q.N(), e = k - from;
        if (e < 1)
            return;
        DenseMatrix Y(r, e);
        for (int j = 0; j < e; j++)
            for (int i = 0; i < r; i++)
                Y(i, j) = st.RV(i, from + j);
        MultiVector W = st.Q * Y;
        MultiVector AW, MW;
        {
            Section s(sections_, "Apply A");
            AW = A_ * W;
        }
        if (mass)
        {
            Section s(sections_, "Apply M");
            MW = M_ * W;
        }
        Section s(sections_, "orthogonalize");
        Blocks blocks(1, &AW);
        Factors factors(1, &st.RA);
        if (mass)
        {
            blocks.push_back(&MW);
            factors.push_back(&st.RM);
        }
        tsqr_append(st, blocks, factors);
    };
    apply_operators(0);

    bool converged_before = false;
    int last_restart = 0;
    const double r0 = B_.norm();
    const long long errors_at_start = BackendHealth<MultiVector>::errors();
    int ret = 1;

    for (int iter = 0; iter < max_iter_; iter++)
    {
        if (BackendHealth<MultiVector>::errors() != errors_at_start)
        {
            std::cerr << "The backend reported a failure; the solve is stopped after "
                      << iter << " iterations" << std::endl;
            return 2;
        }
        const int k = st.RV.N(), r = st.Q.N();
        DenseMatrix const &RM = mass ? st.RM : st.RV;

        DenseMatrix VAV, VBV, VMV;
        {
            Section s(sections_, "Compute VAV");
            VAV = st.RV.transpose() * st.RA;
            DenseMatrix BV = st.RB.transpose() * st.RV;
            VBV = BV.transpose() * BV;
            if (mass)
                VMV = st.RV.transpose() * st.RM;
        }
        if (mass)
            dense_solve_generalized(VAV, VMV, VBV, T);
        else
            dense_solve(VAV, VBV, T);

        // eigenpairs of the residual in the basis: E = R_A T R_M' + R_M T R_A' + R_B R_B'
        DenseMatrix Y(r, r), theta(r, 1);
        std::vector<int> order;
        {
            Section s(sections_, "resid_tsqr");
            DenseMatrix E = st.RA * (T * RM.transpose());
            DenseMatrix E2 = st.RB * st.RB.transpose();
            for (int j = 0; j < r; j++)
                for (int i = 0; i <= j; i++)
                {
                    const double v = E(i, j) + E(j, i) + E2(i, j);
                    E(i, j) = v;
                    if (i != j)
                        E(j, i) = v;
                }
            E.eigs(Y, theta);
            order = detail::largest_by_magnitude(theta, theta.M());
        }
        const double res = theta.M() > 0 ? std::abs(theta(order[0], 0)) : 0.0;

        if (verbose_)
            std::cout << "Iteration " << iter + 1
                      << ". Estimate Lanczos, absolute: " << res
                      << ", relative: " << std::abs(res) / r0 / r0 << std::endl;

        IterationRecord rec;
        rec.iteration = iter + 1;
        rec.space_size = k;
        rec.residual = res;
        rec.relative = std::abs(res) / r0 / r0;
        rec.restarted = false;
        auto record = [&]() {
            rec.seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_begin).count();
            history_.push_back(rec);
            if (observer_)
                observer_(rec);
        };

        const bool converged = std::abs(res) < tol_ * r0 * r0;
        const bool out_of_room = iter + 1 >= max_iter_ || k >= n;
        bool stop = false;
        if (converged || out_of_room)
        {
            if (converged && minimize_solution_space_ && !converged_before)
                converged_before = true;
            else
                stop = true;
        }
        if (stop)
        {
            if (verbose_)
                std::cout << "The Lyapunov solver "
                          << (converged ? "converged" : "did not converge")
                          << " in " << iter + 1
                          << " iterations with a final relative residual of "
                          << res / r0 / r0 << ". The size of the space used "
                          << "for the solution is " << k << std::endl;
            record();
            ret = converged ? 0 : -1;
            break;
        }

        const bool space_full = restart_size_ > 0 && k >= restart_size_;
        const bool period_over = restart_iterations_ > 0 && iter - last_restart >= restart_iterations_;
        const bool at_start = restart_upon_start_ && iter == 0;
        if (space_full || period_over || converged || at_start)
        {
            DenseMatrix X;
            compute_restart_vectors(X, T, std::min(reduced_size_, k), restart_tolerance_);
            if (verbose_)
                std::cout << "Restarted with " << X.N() << " vectors" << std::endl;
            Section s(sections_, "orthogonalize");
            DenseMatrix RV = st.RV * X, RA = st.RA * X;
            detail::assign_in_place(st.RV, RV);
            detail::assign_in_place(st.RA, RA);
            if (mass)
            {
                DenseMatrix RMx = st.RM * X;
                detail::assign_in_place(st.RM, RMx);
            }
            tsqr_shrink(st);
            last_restart = iter;
            rec.restarted = true;
            record();
            continue;
        }

        // expansion: leading residual eigenvectors, orthogonalised against V in
        // coefficient space (V'u = R_V' y); with "Orthogonalize candidates" the
        // nev leading ones are tried and those already inside the space dropped
        // (matlab/RAILSsolver.m:403-409,517-531)
        const int room = (restart_size_ > 0 ? restart_size_ : n) - k;
        const int grow_by = std::min(std::min(expand_size_, theta.M()), room);
        const int nev = orthogonalize_candidates_
                            ? std::min(std::max(lanczos_vectors_ > 0 ? lanczos_vectors_ : expand_size_, grow_by),
                                       theta.M())
                            : grow_by;
        int added = 0;
        for (int c = 0; c < nev && added < grow_by; c++)
        {
            const int col = k + added;
            st.RV.resize(r, col + 1);
            for (int i = 0; i < r; i++)
                st.RV(i, col) = Y(i, order[c]);
            const double nrm = orthogonalize_column(st.RV, col, col, nullptr);
            if (!(nrm > (orthogonalize_candidates_ ? 1e-8 : 0.0)))
            {
                st.RV.resize(r, col);
                continue;
            }
            for (int i = 0; i < r; i++)
                st.RV(i, col) /= nrm;
            added++;
        }
        // R_A / R_M get their new columns (and every block its new rows) here
        apply_operators(k);
        record();
    }

    {
        Section s(sections_, "orthogonalize");
        V = st.Q * st.RV;
    }
    return ret;
}

}
}

#endif
