// Implementation of RAILS::b200::Solver (see RailsSolver.hpp).  Line numbers
// in comments refer to /root/reference/src/LyapunovSolver.hpp.
#ifndef RAILS_B200_RAILSSOLVERIMPL_HPP
#define RAILS_B200_RAILSSOLVERIMPL_HPP

namespace RAILS
{
namespace b200
{

namespace detail
{

// indices of the `count` entries largest in magnitude, in the order the
// reference's find_largest_eigenvalues produces (src/StlTools.hpp:12-30)
template<class DenseMatrix>
std::vector<int> largest_by_magnitude(DenseMatrix const &values, int count)
{
    std::vector<std::pair<int, double> > tagged;
    for (int i = 0; i < values.M(); i++)
        tagged.push_back(std::pair<int, double>(i, values(i)));
    std::sort(tagged.begin(), tagged.end(),
              [](std::pair<int, double> const &a, std::pair<int, double> const &b)
              { return std::abs(a.second) > std::abs(b.second); });
    std::vector<int> picked;
    for (int i = 0; i < count; i++)
        picked.push_back(tagged[i].first);
    return picked;
}

}

// VAV <- [[VAV; W'AV], V'AW],  VBV likewise, AV <- [AV, A W], BV <- [BV, B'W]
// (:141-207).  W holds the columns of V that are new since the last call.
template<class Matrix, class MultiVector, class DenseMatrix>
void Solver<Matrix, MultiVector, DenseMatrix>::project(
    MultiVector const &V, MultiVector const &W, MultiVector &AV, MultiVector &BV, MultiVector &MV,
    DenseMatrix &VAV, DenseMatrix &VBV, DenseMatrix &VMV, bool first, int max_size)
{
    MultiVector AW, BW, MW;
    {
        Section s(sections_, "Apply A");
        AW = A_ * W;
    }
    if (use_mass_)
    {
        // MVnew = M * V(:, new_indices)  (matlab/RAILSsolver.m:368-373)
        Section s(sections_, "Apply M");
        MW = M_ * W;
    }
    {
        Section s(sections_, "Apply B");
        BW = B_.apply_transposed(W);
    }
    if (first)
    {
        // BV has the shape of B', not of V
        BV = MultiVector(BW, max_size);
        BV.resize(0);
    }

    Section s(sections_, "Compute VAV");
    const int old_cols = AV.N();
    const int size = old_cols + W.N();
    VAV.resize(size, size);          // keeps the old entries
    VBV.resize(size, size);

    // the dots of :173-195 (and of the mass-matrix branch), same operands and
    // the same order as the reference; a backend may return them with one host
    // round trip (BatchedDots)
    typedef std::pair<MultiVector const *, MultiVector const *> Pair;
    std::vector<Pair> pairs;
    int i_wav = -1, i_wbv = -1, i_vaw = -1, i_wbw = -1, i_wmv = -1, i_vmw = -1;
    if (old_cols > 0)
    {
        i_wav = (int)pairs.size(); pairs.push_back(Pair(&W, &AV));
        i_wbv = (int)pairs.size(); pairs.push_back(Pair(&BW, &BV));
    }
    i_vaw = (int)pairs.size(); pairs.push_back(Pair(&V, &AW));
    i_wbw = (int)pairs.size(); pairs.push_back(Pair(&BW, &BW));
    if (use_mass_)
    {
        if (old_cols > 0)
        {
            i_wmv = (int)pairs.size(); pairs.push_back(Pair(&W, &MV));
        }
        i_vmw = (int)pairs.size(); pairs.push_back(Pair(&V, &MW));
    }
    std::vector<DenseMatrix> G;
    BatchedDots<MultiVector, DenseMatrix>::run(pairs, G);

    if (old_cols > 0)
    {
        DenseMatrix const &WAV = G[i_wav];
        DenseMatrix const &WBV = G[i_wbv];
        for (int i = 0; i < WAV.M(); i++)
            for (int j = 0; j < WAV.N(); j++)
            {
                VAV(i + old_cols, j) = WAV(i, j);
                VBV(i + old_cols, j) = WBV(i, j);
                VBV(j, i + old_cols) = WBV(i, j);
            }
    }

    DenseMatrix const &VAW = G[i_vaw];
    for (int i = 0; i < VAW.M(); i++)
        for (int j = 0; j < VAW.N(); j++)
            VAV(i, j + old_cols) = VAW(i, j);

    DenseMatrix const &WBW = G[i_wbw];
    for (int i = 0; i < WBW.M(); i++)
        for (int j = 0; j < WBW.N(); j++)
            VBV(i + old_cols, j + old_cols) = WBW(i, j);

    if (use_mass_)
    {
        // VMV = [[VMV; W' MV_old], V' MW]  (matlab/RAILSsolver.m:375-380)
        VMV.resize(size, size);
        if (old_cols > 0)
        {
            DenseMatrix const &WMV = G[i_wmv];
            for (int i = 0; i < WMV.M(); i++)
                for (int j = 0; j < WMV.N(); j++)
                    VMV(i + old_cols, j) = WMV(i, j);
        }
        DenseMatrix const &VMW = G[i_vmw];
        for (int i = 0; i < VMW.M(); i++)
            for (int j = 0; j < VMW.N(); j++)
                VMV(i, j + old_cols) = VMW(i, j);
        MV.push_back(MW);
    }

    AV.push_back(AW);
    BV.push_back(BW);
}

// Restart: V <- V X, AV <- AV X, BV <- BV X, VAV <- X'VAV X, VBV <- X'VBV X
// with X the dominant eigenvectors of T (:258-303)
template<class Matrix, class MultiVector, class DenseMatrix>
void Solver<Matrix, MultiVector, DenseMatrix>::compress(
    MultiVector &V, MultiVector &W, MultiVector &AV, MultiVector &BV, MultiVector &MV,
    DenseMatrix &VAV, DenseMatrix &VBV, DenseMatrix &VMV, DenseMatrix const &T, TsqrState *tsqr)
{
    DenseMatrix X;
    compute_restart_vectors(X, T, std::min(reduced_size_, V.N()), restart_tolerance_);
    const int kept = X.N();
    Section s(sections_, "Restart compression");

    if (tsqr)
    {
        // V X and AV X are combinations of columns the basis already spans:
        // the factorisation follows with two small products, Q is untouched
        DenseMatrix RV = tsqr->RV * X, RA = tsqr->RA * X;
        tsqr->RV = RV;
        tsqr->RA = RA;
        tsqr->columns = kept;
        // ... unless the basis has grown well beyond what the compressed factor
        // needs: then it is rebuilt from [B, V, AV] at the next estimate
        const int needed = 2 * kept + (B_.is_multivector() ? B_.multivector().N() : 0);
        if (tsqr->Q.N() > 2 * needed + 8)
            tsqr->valid = false;
    }

    V.view(0, kept - 1) = V * X;
    V.resize(kept);

    if (verbose_)
        std::cout << "Restarted with " << V.N() << " vectors" << std::endl;

    W.resize(0);

    DenseMatrix small = X.transpose() * (VAV * X);
    VAV.resize(kept, kept);
    VAV.view() = small;

    AV.view(0, kept - 1) = AV * X;
    AV.resize(kept);

    small = X.transpose() * (VBV * X);
    VBV.resize(kept, kept);
    VBV.view() = small;

    BV.view(0, kept - 1) = BV * X;
    BV.resize(kept);

    if (use_mass_)
    {
        small = X.transpose() * (VMV * X);
        VMV.resize(kept, kept);
        VMV.view() = small;

        MV.view(0, kept - 1) = MV * X;
        MV.resize(kept);
    }
}

template<class Matrix, class MultiVector, class DenseMatrix>
int Solver<Matrix, MultiVector, DenseMatrix>::solve(MultiVector &V, DenseMatrix &T)
{
    const auto t_begin = std::chrono::steady_clock::now();
    const bool trace = std::getenv("RAILS_TRACE") != nullptr;
    std::map<std::string, double> traced;
    auto trace_iteration = [&](int it) {
        if (!trace) return;
        std::cerr << "trace " << it << ":";
        for (auto const &kv : sections_)
        {
            std::cerr << " " << kv.first << "=" << 1e3 * (kv.second - traced[kv.first]);
            traced[kv.first] = kv.second;
        }
        std::cerr << std::endl;
    };
    // opts.projection_method (matlab/RAILSsolver.m:7-17): which of the A^-1 variants
    const double pm_fraction = projection_method_ - std::floor(projection_method_);
    const bool pm_inverse = projection_method_ > 1.0 + 1e-8;
    const bool pm_extended = pm_inverse && std::floor(projection_method_) == 2.0;   // [r, A^-1 r]
    const bool pm_start_1 = std::abs(pm_fraction - 0.1) < 1e-8;
    const bool pm_start_2 = std::abs(pm_fraction - 0.2) < 1e-8;
    const bool pm_start_3 = std::abs(pm_fraction - 0.3) < 1e-8;
    if (pm_inverse && !has_inverse_)
    {
        std::cerr << "Projection method " << projection_method_
                  << " needs an inverse operator (set_inverse)" << std::endl;
        return 1;
    }
    if (!pm_inverse && has_inverse_ && verbose_)
        std::cout << "An inverse application method is provided, but the current "
                  << "projection method does not make use of this" << std::endl;   // :280-283
    if (pm_start_2 && !B_.is_multivector())
    {
        std::cerr << "Projection method " << projection_method_
                  << " starts from B, which must be a MultiVector" << std::endl;
        return 1;
    }

    if (residual_method_ == 2 && B_.is_multivector() && !m_orthogonalization_ && !has_nullspace_ &&
        !pm_inverse)
        return solve_coefficient_space(V, T);

    history_.clear();
    sections_.clear();

    const int n = V.M();
    int max_size = std::max(V.N(), std::min(restart_size_ > 0 ? restart_size_ : 100, n));
    const bool custom_ortho = m_orthogonalization_ || has_nullspace_ || pm_inverse;

    // W -> A^-1 W (1.x) or [W, A^-1 W] (2.x)  (matlab/RAILSsolver.m:520-524)
    auto with_inverse = [&](MultiVector const &block) {
        if (!pm_inverse || block.N() < 1)
            return block.copy();
        Section s(sections_, "Apply inverse");
        MultiVector inv = Ainv_ * block;
        if (!pm_extended)
            return inv;
        MultiVector both(block, 2 * block.N());
        both.resize(0);
        both.push_back(block);
        both.push_back(inv);
        return both;
    };

    if (!restart_from_solution_)
    {
        // random normalised start vector in a buffer with room to grow (:108-115)
        V.resize(max_size);
        V.resize(1);
        V.random();
        // with an inverse projection method the start space is orthonormalised
        // further down, after A^-1 has been applied
        if (custom_ortho && !pm_inverse)
        {
            // V = Morth(V, M, nullspace) (matlab/RAILSsolver.m:303-307)
            MultiVector start = V.copy();
            V.resize(0);
            morth(start, has_nullspace_ ? &nullspace_ : (MultiVector const *) nullptr, V);
        }
        else if (!pm_inverse)
            V.orthogonalize();
    }
    else if (max_size != V.N())
    {
        const int have = V.N();
        V.resize(max_size);
        V.resize(have);
    }

    if (pm_inverse && (pm_start_1 || pm_start_2 || !restart_from_solution_))
    {
        // matlab/RAILSsolver.m:290-307: V0 = the given or random space (x.1, x.3) or
        // B (x.2); W = A^-1 V0 (x.1, x.2); V = [V0, W] (2.1, 2.2), W (1.1, 1.2), V0 (x.3)
        MultiVector start = pm_start_2 ? B_.multivector().copy() : V.copy();
        if (pm_start_1 || pm_start_2)
            start = with_inverse(start);
        else if (!pm_start_3)
            std::cerr << "Projection method " << projection_method_
                      << ": unknown start variant, starting from V_0" << std::endl;
        if (start.N() > max_size)
        {
            max_size = start.N();
            V.resize(max_size);
        }
        V.resize(0);
        morth(start, has_nullspace_ ? &nullspace_ : (MultiVector const *) nullptr, V);
    }

    MultiVector W(V);                 // deep copy, as `MultiVector W = V` (:123)

    MultiVector AV(V, max_size);
    DenseMatrix VAV(max_size, max_size);
    AV.resize(0);

    MultiVector BV;
    DenseMatrix VBV(max_size, max_size);

    MultiVector MV;
    DenseMatrix VMV;
    if (use_mass_)
    {
        MV = MultiVector(V, max_size);
        MV.resize(0);
        VMV = DenseMatrix(max_size, max_size);
    }

    TsqrState tsqr;                  // "Residual method" = 1: QR of [V, AV, B] kept across iterations
    bool converged_before = false;
    int last_restart = 0;
    const double r0 = B_.norm();
    const long long errors_at_start = BackendHealth<MultiVector>::errors();

    for (int iter = 0; iter < max_iter_; iter++)
    {
        if (BackendHealth<MultiVector>::errors() != errors_at_start)
        {
            std::cerr << "The backend reported a failure; the solve is stopped after "
                      << iter << " iterations" << std::endl;
            return 2;
        }
        const int cols_before = V.N();
        if (W.N())
            project(V, W, AV, BV, MV, VAV, VBV, VMV, iter == 0, max_size);

        // the Lanczos start vector is drawn here rather than after the projected
        // solve (neither draws from std::rand() in between, so the random stream
        // is the reference's): the GPU projects it while the host works on T
        MultiVector Qstart;
        const bool started = residual_method_ == 0 &&
                             lanczos_start(AV, use_mass_ ? MV : V, Qstart, lanczos_iterations_);

        // an M-orthonormal basis has V'MV = I: the standard projected equation
        // (matlab/RAILSsolver.m:375-385, `hasM && ~mortho`)
        if (use_mass_ && !m_orthogonalization_)
            dense_solve_generalized(VAV, VMV, VBV, T);
        else
            dense_solve(VAV, VBV, T);

        const int L = lanczos_iterations_;
        DenseMatrix H(L + 1, L + 1);
        DenseMatrix ritz_values(L, 1);
        MultiVector ritz_vectors;
        // generalised residual AV T MV' + MV T AV' + B B' (matlab/RAILSsolver.m:391)
        bool have_ritz = false;
        if (residual_method_ >= 1 && B_.is_multivector())
        {
            const int want = std::max(expand_size_, orthogonalize_candidates_ ? lanczos_vectors_ : 0);
            have_ritz = resid_tsqr(tsqr, AV, use_mass_ ? MV : V, T, ritz_vectors, ritz_values, want) == 0;
        }
        if (!have_ritz)
            resid_lanczos(AV, use_mass_ ? MV : V, T, H, ritz_vectors, ritz_values, L,
                          started ? &Qstart : (MultiVector *) nullptr);

        // the QR path returns its eigenvalues sorted and only the leading
        // eigenvectors: pick those in order (a tie must not send the unstable
        // sort of the reference's selection to a column that was not computed)
        auto pick_ritz = [&](int count) {
            if (!have_ritz)
                return detail::largest_by_magnitude(ritz_values, count);
            std::vector<int> first(count);
            for (int i = 0; i < count; i++)
                first[i] = i;
            return first;
        };

        const double res = ritz_values.norm_inf();

        if (verbose_)
            std::cout << "Iteration " << iter + 1
                      << ". Estimate Lanczos, absolute: " << res
                      << ", relative: " << std::abs(res) / r0 / r0 << std::endl;

        IterationRecord rec;
        rec.iteration = iter + 1;
        rec.space_size = V.N();
        rec.residual = res;
        rec.relative = std::abs(res) / r0 / r0;
        rec.restarted = false;

        const bool converged = std::abs(res) < tol_ * r0 * r0;
        const bool out_of_room = iter + 1 >= max_iter_ || V.N() >= n;
        bool stop = false;
        if (converged || out_of_room)
        {
            // first convergence only triggers a compression restart (:226-229)
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
                          << "for the solution is " << V.N() << std::endl;
            rec.seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_begin).count();
            trace_iteration(iter + 1);
            history_.push_back(rec);
            if (observer_)
                observer_(rec);
            return converged ? 0 : -1;
        }

        const bool space_full = restart_size_ > 0 && V.N() >= restart_size_;
        const bool period_over = restart_iterations_ > 0 && iter - last_restart >= restart_iterations_;
        const bool at_start = restart_upon_start_ && iter == 0;    // matlab/RAILSsolver.m:456
        if (space_full || period_over || converged || at_start)
        {
            if (verbose_)
            {
                if (converged)
                    std::cout << "Method converged. Minimizing the solution "
                              << "space size";
                else if (restart_size_ > 0)
                    std::cout << "Reached the maximum space size of " << restart_size_;
                else if (at_start)
                    std::cout << "Restart upon start";
                else if (restart_iterations_ > 0)
                    std::cout << restart_iterations_ << " iterations have passed";
                else
                    std::cout << "No clue what happened";
                std::cout << ". Trying to restart with "
                          << (reduced_size_ > 0 ? reduced_size_ : V.N())
                          << " vectors" << std::endl;
            }

            compress(V, W, AV, BV, MV, VAV, VBV, VMV, T, tsqr.valid ? &tsqr : (TsqrState *) nullptr);
            last_restart = iter;

            rec.restarted = true;
            rec.seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_begin).count();
            trace_iteration(iter + 1);
            history_.push_back(rec);
            if (observer_)
                observer_(rec);
            continue;
        }

        const int room = (restart_size_ > 0 ? restart_size_ : n) - V.N();
        const int grow_by = std::min(std::min(expand_size_, ritz_values.M()), room);

        // capacity grows 100 columns at a time (:311-332)
        const int add_cols = pm_extended ? 2 * grow_by : grow_by;
        if (V.N() + add_cols > max_size)
        {
            Section s(sections_, "Resize spaces");
            max_size += std::max(100, add_cols);
            const int have = V.N();

            V.resize(max_size);
            V.resize(have);

            AV.resize(max_size);
            AV.resize(have);

            VAV.resize(max_size, max_size);
            VAV.resize(have, have);

            BV.resize(max_size);
            BV.resize(have);

            VBV.resize(max_size, max_size);
            VBV.resize(have, have);

            if (use_mass_)
            {
                MV.resize(max_size);
                MV.resize(have);

                VMV.resize(max_size, max_size);
                VMV.resize(have, have);
            }
        }

        // expand with the Ritz vectors of the largest residual Ritz values
        int appended = grow_by;
        if (orthogonalize_candidates_)
        {
            // matlab/RAILSsolver.m:403-409,517-531: the nev leading candidates are
            // orthonormalised against V first (those already inside the space
            // drop out), the first `grow_by` survivors are then appended
            Section s(sections_, "orthogonalize");
            const int nev = std::min(lanczos_vectors_ > 0 ? lanczos_vectors_ : expand_size_,
                                     ritz_values.M());
            std::vector<int> picked = pick_ritz(nev);
            MultiVector candidates(V, std::max(nev, 1));
            candidates.resize(0);
            for (int i = 0; i < nev; i++)
                candidates.push_back(ritz_vectors.view(picked[i]));
            MultiVector survivors(V, std::max(nev, 1));
            survivors.resize(0);
            morth(candidates, &V, survivors);
            survivors.resize(std::min(grow_by, survivors.N()));
            MultiVector block = with_inverse(survivors);
            appended = morth(block, has_nullspace_ ? &nullspace_ : (MultiVector const *) nullptr, V);
        }
        else if (custom_ortho)
        {
            Section s(sections_, "orthogonalize");
            std::vector<int> picked = pick_ritz(grow_by);
            MultiVector block(V, std::max(grow_by, 1));
            block.resize(0);
            for (int i = 0; i < grow_by; i++)
                block.push_back(ritz_vectors.view(picked[i]));
            MultiVector expansion = with_inverse(block);
            appended = morth(expansion, has_nullspace_ ? &nullspace_ : (MultiVector const *) nullptr, V);
        }
        else
        {
            {
                Section e(sections_, "Expand");
                std::vector<int> picked = pick_ritz(grow_by);
                for (int i = 0; i < grow_by; i++)
                    V.push_back(ritz_vectors.view(picked[i]));
            }
            Section s(sections_, "orthogonalize");
            V.orthogonalize();
        }

        if (appended > 0)
            W = V.view(cols_before, cols_before + appended - 1);
        else
            W.resize(0);      // every candidate was already in the space

        rec.seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_begin).count();
        trace_iteration(iter + 1);
        history_.push_back(rec);
        if (observer_)
            observer_(rec);
    }

    return 1;
}

template<class Matrix, class MultiVector, class DenseMatrix>
int Solver<Matrix, MultiVector, DenseMatrix>::dense_solve(DenseMatrix const &A,
                                                          DenseMatrix const &B, DenseMatrix &X)
{
    Section s(sections_, "dense_solve");
    // sb03md('C','X','N','T'): A Y + Y A' = scale * C, solution returned in
    // the C array; then X = -Y (:348-365, src/SlicotWrapper.cpp:8-49)
    X = B.copy();
    DenseMatrix S = A.copy();
    int n = A.M();
    int info = 0;
    double scale = 1.0;
    if (n < 1)
    {
        std::cerr << "Slicot error: n < 1 is not supported" << std::endl;
        return info;
    }

    char dico = 'C', job = 'X', fact = 'N', trans = 'T';
    const int pad = n + 2;                     // workspace sizes of the reference wrapper
    std::vector<double> U((size_t)pad * pad), wr(pad), wi(pad);
    int ldwork = 2 * std::max(pad * pad, 3 * pad);
    std::vector<double> work(ldwork);
    double sep = 0.0, ferr = 0.0;
    int lda = S.LDA(), ldx = X.LDA();
    sb03md_(&dico, &job, &fact, &trans, &n, S, &lda, U.data(), &n, X, &ldx, &scale, &sep,
            &ferr, wr.data(), wi.data(), NULL, work.data(), &ldwork, &info);

    X *= -1.0;

    if (info != 0 && info != n + 1)
        std::cerr << "Error: sb03md returned info = " << info << std::endl;
    return info;
}

// A X E' + E X A' + B = 0 with E = V'MV (symmetric positive definite for an
// SPD mass matrix and a full-rank V), reduced to the standard problem
//     (E^-1 A) X + X (E^-1 A)' + E^-1 B E^-T = 0
// and handed to dense_solve.  MATLAB calls lyap(A, B, [], E)
// (matlab/RAILSsolver.m:381 -> matlab/mex/lyap.c:55-84, SLICOT sg03ad).
template<class Matrix, class MultiVector, class DenseMatrix>
int Solver<Matrix, MultiVector, DenseMatrix>::dense_solve_generalized(
    DenseMatrix const &A, DenseMatrix const &E, DenseMatrix const &B, DenseMatrix &X)
{
    int n = A.M();
    if (n < 1)
        return dense_solve(A, B, X);
    DenseMatrix Ah, Bh;
    {
        Section s(sections_, "dense_solve");
        DenseMatrix lu = E.copy();
        Ah = A.copy();
        DenseMatrix Y = B.copy();
        std::vector<int> piv(n);
        int info = 0, lda = lu.LDA(), ldb = Ah.LDA();
        // one factorisation, two right-hand-side blocks
        dgesv_(&n, &n, lu, &lda, piv.data(), Ah, &ldb, &info);
        if (info)
        {
            std::cerr << "Error: V'MV is singular (dgesv info = " << info << ")" << std::endl;
            return info;
        }
        // Y = E^-1 B, then Bh = Y E^-T = (E^-1 Y')'
        DenseMatrix lu2 = E.copy();
        int ldy = Y.LDA();
        lda = lu2.LDA();
        dgesv_(&n, &n, lu2, &lda, piv.data(), Y, &ldy, &info);
        DenseMatrix Yt(n, n);
        for (int j = 0; j < n; j++)
            for (int i = 0; i < n; i++)
                Yt(i, j) = Y(j, i);
        DenseMatrix lu3 = E.copy();
        int ldt = Yt.LDA();
        lda = lu3.LDA();
        dgesv_(&n, &n, lu3, &lda, piv.data(), Yt, &ldt, &info);
        Bh = DenseMatrix(n, n);
        for (int j = 0; j < n; j++)
            for (int i = 0; i < n; i++)
                Bh(i, j) = Yt(j, i);
    }
    return dense_solve(Ah, Bh, X);
}

// L steps of symmetric Lanczos (no re-orthogonalisation) on the residual
// operator applied through its low-rank factors (:367-447):
//     R q = B (B' q) + AV (T (V' q)) + V (T (AV' q))
template<class Matrix, class MultiVector, class DenseMatrix>
bool Solver<Matrix, MultiVector, DenseMatrix>::lanczos_start(
    MultiVector const &AV, MultiVector const &V, MultiVector &Q, int max_iter)
{
    if (!(FusedLanczos<MultiVector, DenseMatrix>::enabled && B_.is_multivector() &&
          FusedLanczos<MultiVector, DenseMatrix>::active() &&
          FusedLanczos<MultiVector, DenseMatrix>::fits(V.N(), B_.multivector().N())))
        return false;
    Section s(sections_, "resid_lanczos");
    Q = MultiVector(V, max_iter + 1);
    Q.resize(1);
    Q.random();
    FusedLanczos<MultiVector, DenseMatrix>::begin(AV, V, B_.multivector(), Q, max_iter);
    return true;
}

template<class Matrix, class MultiVector, class DenseMatrix>
int Solver<Matrix, MultiVector, DenseMatrix>::resid_lanczos(
    MultiVector const &AV, MultiVector const &V, DenseMatrix const &T, DenseMatrix &H,
    MultiVector &eigenvectors, DenseMatrix &eigenvalues, int max_iter)
{
    return resid_lanczos(AV, V, T, H, eigenvectors, eigenvalues, max_iter, (MultiVector *) nullptr);
}

template<class Matrix, class MultiVector, class DenseMatrix>
int Solver<Matrix, MultiVector, DenseMatrix>::resid_lanczos(
    MultiVector const &AV, MultiVector const &V, DenseMatrix const &T, DenseMatrix &H,
    MultiVector &eigenvectors, DenseMatrix &eigenvalues, int max_iter, MultiVector *started)
{
    Section s(sections_, "resid_lanczos");

    MultiVector Q;
    if (started)
        Q = *started;              // shallow: shares the buffer the backend is working on
    else
    {
        Q = MultiVector(V, max_iter + 1);
        Q.resize(1);
        Q.random();
    }

    H = 0.0;

    int steps = 0;
    if (FusedLanczos<MultiVector, DenseMatrix>::enabled && B_.is_multivector() &&
        FusedLanczos<MultiVector, DenseMatrix>::active() &&
        FusedLanczos<MultiVector, DenseMatrix>::fits(V.N(), B_.multivector().N()))
    {
        // same recurrence, three fused kernels per step (backend-supplied)
        steps = FusedLanczos<MultiVector, DenseMatrix>::run(AV, V, T, B_.multivector(), H, Q,
                                                            max_iter);
    }
    else
    {
    Q.view(0) /= Q.norm();

    double alpha = 0.0, beta = 0.0;
    for (int i = 0; i < max_iter; i++)
    {
        Q.resize(steps + 2);
        // bound to the view temporaries themselves: copy-constructing a view
        // would deep-copy it (the concept's copy constructor is deep)
        MultiVector const &q = Q.view(steps);
        MultiVector &&next = Q.view(steps + 1);

        MultiVector Y = B_.apply_transposed(q);
        next = B_.apply(Y);

        DenseMatrix Z = V.dot(q);
        Z = T * Z;
        next += AV * Z;

        Z = AV.dot(q);
        Z = T * Z;
        next += V * Z;

        alpha = next.dot(q)(0, 0);
        H(steps, steps) = alpha;

        next -= alpha * q;
        if (steps > 0)
            next -= beta * Q.view(steps - 1);

        beta = next.norm();
        if (std::getenv("RAILS_DEBUG_LANCZOS"))
            std::cerr << "lanczos plain step " << steps << " alpha " << alpha << " beta " << beta
                      << std::endl;
        if (beta < 1e-14)
        {
            // invariant subspace: keep H as it is, cannot normalise
            steps++;
            break;
        }

        H(steps + 1, steps) = beta;
        H(steps, steps + 1) = beta;

        next /= beta;
        steps++;
    }
    }

    H.resize(steps, steps);
    Q.resize(steps);

    DenseMatrix v(steps, steps);
    H.eigs(v, eigenvalues);

    eigenvectors = Q * v;
    return 0;
}

// Orthogonalises the columns of the given blocks against the basis kept in
// `state`, appends the independent ones to the basis and the coefficient
// columns to the R factor of each block: afterwards blocks[i] = Q * Rs[i][:, new
// columns] to rounding.  All columns are projected against the OLD basis
// together (two classical Gram-Schmidt passes = two Q'W and two W -= Q C
// products, whatever the number of columns: Q is streamed four times per call),
// then one by one against the few columns this call has added.  A column whose
// remainder is below 1e-10 of its norm is dependent -- from the second iteration
// on that is every expansion vector, an eigenvector of the previous residual and
// therefore a combination of the columns already in F -- and only contributes
// its coefficients.  Dropping it is what makes two passes enough: normalising
// such a remainder would bring back O(1) components along the old columns.
template<class Matrix, class MultiVector, class DenseMatrix>
void Solver<Matrix, MultiVector, DenseMatrix>::tsqr_append(
    TsqrState &state, std::vector<MultiVector const *> const &blocks,
    std::vector<DenseMatrix *> const &Rs)
{
    int total = 0;
    for (size_t b = 0; b < blocks.size(); b++)
        total += blocks[b]->N();
    if (total < 1)
        return;

    MultiVector W(*blocks[0], total);
    W.resize(0);
    for (size_t b = 0; b < blocks.size(); b++)
        W.push_back(*blocks[b]);

    std::vector<double> full(total);
    for (int j = 0; j < total; j++)
    {
        MultiVector const &w = static_cast<MultiVector const &>(W).view(j);
        full[j] = w.norm();
    }

    const int r0 = state.Q.N();
    DenseMatrix coef(std::max(r0, 1), total);
    coef.resize(r0, total);
    coef = 0.0;
    for (int pass = 0; pass < 2 && r0 > 0; pass++)
    {
        DenseMatrix C = state.Q.dot(W);
        W -= state.Q * C;
        for (int j = 0; j < total; j++)
            for (int i = 0; i < r0; i++)
                coef(i, j) += C(i, j);
    }

    if (r0 + total > state.capacity)
    {
        state.capacity = std::min(W.M(), std::max(state.capacity + 100, r0 + total));
        state.Q.resize(state.capacity);
        state.Q.resize(r0);
    }

    int col = 0;
    for (size_t b = 0; b < blocks.size(); b++)
        for (int jb = 0; jb < blocks[b]->N(); jb++, col++)
        {
            MultiVector &&w = W.view(col);
            const int r = state.Q.N();                 // r0 + columns added by this call
            std::vector<double> extra(r - r0, 0.0);
            if (r > r0)
            {
                MultiVector const &added = static_cast<MultiVector const &>(state.Q).view(r0, r - 1);
                for (int pass = 0; pass < 2; pass++)
                {
                    DenseMatrix c = added.dot(w);
                    w -= added * c;
                    for (int i = 0; i < r - r0; i++)
                        extra[i] += c(i, 0);
                }
            }
            const double rest = w.norm();
            const bool independent = full[col] > 0.0 && rest > 1e-10 * full[col] &&
                                     state.Q.N() < state.capacity;
            if (independent)
            {
                w /= rest;
                state.Q.push_back(w);
            }
            // every R block carries one row per basis vector
            const int rows = state.Q.N();
            DenseMatrix *all[4] = {&state.RV, &state.RA, &state.RB, &state.RM};
            for (int q = 0; q < 4; q++)
            {
                DenseMatrix &X = *all[q];
                const int have_rows = X.M(), have_cols = X.N();
                if (have_rows == rows)
                    continue;
                X.resize(rows, have_cols);
                for (int c2 = 0; c2 < have_cols; c2++)
                    for (int i = have_rows; i < rows; i++)
                        X(i, c2) = 0.0;
            }
            DenseMatrix &R = *Rs[b];
            const int at = R.N();
            R.resize(rows, at + 1);
            for (int i = 0; i < rows; i++)
                R(i, at) = i < r0 ? coef(i, col) : (i < r ? extra[i - r0] : (independent ? rest : 0.0));
        }
}

// Non-zero eigenpairs of the residual R = F S F' from a QR factorisation of its
// tall-skinny factor (see the declaration), kept up to date incrementally:
//     F = [V, AV, B] = Q [R_V, R_AV, R_B]
//     E = R_V T R_AV' + R_AV T R_V' + R_B R_B'          (r x r, r = rank F)
//     E = Y Theta Y'   ->   eigenvectors Q Y
// Each outer iteration only orthogonalises the columns that are new since the
// last call (e columns of V and of AV, as one block): 4 passes over Q instead of
// the 2 L passes over [V, AV, B] of the Lanczos estimate, and one Q * Y product for the
// `want` leading eigenvectors only.  eigenvalues come back sorted by decreasing
// magnitude (all r of them), eigenvectors holds the first `want`.
template<class Matrix, class MultiVector, class DenseMatrix>
int Solver<Matrix, MultiVector, DenseMatrix>::resid_tsqr(
    TsqrState &state, MultiVector const &AV, MultiVector const &V, DenseMatrix const &T,
    MultiVector &eigenvectors, DenseMatrix &eigenvalues, int want)
{
    Section s(sections_, "resid_tsqr");
    MultiVector const &B = B_.multivector();
    const int k = V.N(), nb = B.N();

    if (!state.valid)
    {
        state.capacity = std::min(V.M(), std::max(100, 2 * k + nb + 2));
        state.Q = MultiVector(V, state.capacity);
        state.Q.resize(0);
        // capacity up front: resize() keeps entries in place within it
        const int kcap = std::max(V.N(), std::min(restart_size_ > 0 ? restart_size_ : 100, V.M())) + 1;
        state.RV = DenseMatrix(state.capacity, kcap); state.RV.resize(0, 0);
        state.RA = DenseMatrix(state.capacity, kcap); state.RA.resize(0, 0);
        state.RB = DenseMatrix(state.capacity, std::max(nb, 1)); state.RB.resize(0, 0);
        state.RM = DenseMatrix(1, 1); state.RM.resize(0, 0);
        state.columns = 0;
        MultiVector const &Ball = B.view(0, nb - 1);
        tsqr_append(state, std::vector<MultiVector const *>(1, &Ball),
                    std::vector<DenseMatrix *>(1, &state.RB));
        state.valid = true;
    }
    if (state.columns < k)
    {
        // bound to the view temporaries (a copy-constructed view would be a deep copy)
        MultiVector const &Vnew = V.view(state.columns, k - 1);
        MultiVector const &AVnew = AV.view(state.columns, k - 1);
        std::vector<MultiVector const *> blocks;
        std::vector<DenseMatrix *> factors;
        blocks.push_back(&Vnew); factors.push_back(&state.RV);
        blocks.push_back(&AVnew); factors.push_back(&state.RA);
        tsqr_append(state, blocks, factors);
        state.columns = k;
    }

    const int r = state.Q.N();
    DenseMatrix E = state.RV * (T * state.RA.transpose());
    DenseMatrix E2 = state.RB * state.RB.transpose();
    for (int j = 0; j < r; j++)
        for (int i = 0; i <= j; i++)
        {
            const double v = E(i, j) + E(j, i) + E2(i, j);
            E(i, j) = v;
            if (i != j)
                E(j, i) = v;
        }
    DenseMatrix Y(r, r), theta(r, 1);
    E.eigs(Y, theta);
    for (int i = 0; i < theta.M(); i++)
        if (!(theta(i, 0) == theta(i, 0)))
            return 1;

    const int found = theta.M();
    std::vector<int> order = detail::largest_by_magnitude(theta, found);
    want = std::max(1, std::min(want, found));
    eigenvalues = DenseMatrix(found, 1);
    DenseMatrix C(r, want);
    for (int c = 0; c < found; c++)
    {
        eigenvalues(c, 0) = theta(order[c], 0);
        if (c < want)
            for (int i = 0; i < r; i++)
                C(i, c) = Y(i, order[c]);
    }
    eigenvectors = state.Q * C;
    return 0;
}

template<class Matrix, class MultiVector, class DenseMatrix>
void Solver<Matrix, MultiVector, DenseMatrix>::set_nullspace(MultiVector const &N)
{
    nullspace_ = N.copy();
    nullspace_.orthogonalize();       // Morth(opts.nullspace, []) (matlab/RAILSsolver.m:218-220)
    has_nullspace_ = nullspace_.N() > 0;
}

// Morth (matlab/RAILSsolver.m:538-617), see the declaration
template<class Matrix, class MultiVector, class DenseMatrix>
int Solver<Matrix, MultiVector, DenseMatrix>::morth(
    MultiVector const &W, MultiVector const *against, MultiVector &V)
{
    const bool m_inner = m_orthogonalization_;
    // w <- w - Q (Q' w), or w - Q (Q' M w) in the M inner product (:584-592,606-613)
    auto remove = [&](MultiVector &w, MultiVector const &Q) {
        if (Q.N() < 1)
            return;
        DenseMatrix c = m_inner ? Q.dot(M_ * w) : Q.dot(w);
        w -= Q * c;
    };
    int appended = 0;
    for (int i = 0; i < W.N(); i++)
    {
        MultiVector w = W.view(i).copy();
        double nrm = w.norm();                  // v1 = W(:,i) / norm(W(:,i))
        if (!(nrm > 0.0))
            continue;
        w /= nrm;
        for (int pass = 0; pass < 2; pass++)
        {
            if (against)
                remove(w, *against);
            remove(w, V);
        }
        if (m_inner)
        {
            DenseMatrix g = w.dot(M_ * w);      // nrm = sqrt(v1' M v1)
            nrm = g(0, 0) > 0.0 ? std::sqrt(g(0, 0)) : 0.0;
        }
        else
            nrm = w.norm();
        if (!(nrm >= 1e-8))                     // tol = 1e-8 (:554,578-580)
            continue;
        w /= nrm;
        V.push_back(w);
        appended++;
    }
    return appended;
}

template<class Matrix, class MultiVector, class DenseMatrix>
int Solver<Matrix, MultiVector, DenseMatrix>::compute_restart_vectors(
    DenseMatrix &X, DenseMatrix const &T, int num, double tol)
{
    Section s(sections_, "compute_restart_vectors");

    // all eigenpairs of T (DSYEV 'V','U' with a workspace query, as
    // src/LapackWrapper.cpp:18-39)
    DenseMatrix vecs = T.copy();
    DenseMatrix vals(T.N(), 1);
    {
        char jobz = 'V', uplo = 'U';
        int n = vecs.N(), lda = vecs.LDA(), info = 0, lwork = -1;
        double query = 0.0;
        dsyev_(&jobz, &uplo, &n, vecs, &lda, vals, &query, &lwork, &info);
        lwork = (int)query;
        std::vector<double> work(std::max(lwork, 1));
        dsyev_(&jobz, &uplo, &n, vecs, &lda, vals, work.data(), &lwork, &info);
        if (info)
            std::cerr << "DSYEV info = " << info << std::endl;
    }

    num = (num > 0 ? num : T.N());
    X = DenseMatrix(T.N(), num);

    std::vector<int> picked = detail::largest_by_magnitude(vals, num);

    // candidates are sorted by magnitude, so the rejected ones are a suffix
    // and column i can be written in place (:468-479)
    // MATLAB cuts relative to the largest eigenvalue (matlab/RAILSsolver.m:476-477),
    // the C++ reference on the absolute value (:468-479)
    double scale = 1.0;
    if (relative_restart_tolerance_ && num > 0)
        scale = std::max(std::abs(vals(picked[0], 0)), 1e-300);
    int kept = 0;
    for (int i = 0; i < num; ++i)
        if (std::abs(vals(picked[i], 0)) / scale > tol)
        {
            for (int j = 0; j < T.N(); ++j)
                X(j, i) = vecs(j, picked[i]);
            kept++;
        }

    X.resize(T.N(), kept);
    return 0;
}

}
}

#endif
