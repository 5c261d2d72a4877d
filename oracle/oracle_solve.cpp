// TEST INFRASTRUCTURE -- NOT PART OF THE PRODUCT (see oracle_kernels.cpp).
//
// oracle_solve: the outer RAILS loop (/root/reference/src/LyapunovSolver.hpp:
// 100-346) restated on flat arrays.  Same entry-point shape as
// ref_stl_solve / rails_host_solve so one Python harness drives all arms.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <map>
#include <string>
#include <vector>

extern "C" {
void oracle_spmm(int m, const int *rowptr, const int *colidx, const double *vals, int ncols,
                 const double *X, long long ldx, double *Y, long long ldy);
void oracle_dot(long long m, int a, const double *X, long long ldx, int b, const double *Y,
                long long ldy, double *out);
void oracle_gemm(long long m, int k, const double *V, long long ldv, const double *C, int ldc,
                 int c, double alpha, double beta, double *Out, long long ldo);
void oracle_orthogonalize(long long m, int n, double *V, long long ld, int first);
void oracle_random(long long m, int n, unsigned seed, long long row_begin, long long row_end,
                   double *X, long long ld);
int oracle_dense_solve(int k, const double *A, int lda, const double *C, int ldc, double *X, int ldx);
int oracle_resid_lanczos(long long m, int k, const double *V, long long ldv, const double *AV,
                         long long ldav, const double *T, int ldt, int n_b, const double *B,
                         long long ldb, int L, unsigned seed, double *Q, long long ldq,
                         double *H, int ldh);
void dsyev_(const char *jobz, const char *uplo, const int *n, double *a, const int *lda,
            double *w, double *work, const int *lwork, int *info);
void dpotrf_(const char *uplo, const int *n, double *a, const int *lda, int *info);
void dtrsm_(const char *side, const char *uplo, const char *transa, const char *diag, const int *m,
            const int *n, const double *alpha, const double *a, const int *lda, double *b,
            const int *ldb);
}

namespace
{

typedef std::vector<double> vec;

struct Params
{
    std::map<std::string, double> kv;
    // as given / UPPER / lower / Title Case, last present wins (:40-70)
    double get(std::string const &name, double def) const
    {
        std::string up(name), lo(name), ti(name);
        for (size_t i = 0; i < name.size(); ++i)
        {
            up[i] = (char)std::toupper(name[i]);
            lo[i] = (char)std::tolower(name[i]);
        }
        if (!ti.empty()) ti[0] = (char)std::toupper(ti[0]);
        for (size_t i = 1; i < ti.size(); ++i)
            if (!isalpha(ti[i - 1]) && islower(ti[i])) ti[i] = (char)std::toupper(ti[i]);
        double v = def;
        const std::string *order[4] = {&name, &up, &lo, &ti};
        for (auto *s : order)
        {
            auto it = kv.find(*s);
            if (it != kv.end()) v = it->second;
        }
        return v;
    }
};

int sym_eig(int n, double *a, int lda, double *w)
{
    if (n == 0) return 0;
    int info = 0, lwork = -1;
    double q = 0.0;
    dsyev_("V", "U", &n, a, &lda, w, &q, &lwork, &info);
    lwork = std::max(3 * n, (int)q);
    vec work(lwork);
    dsyev_("V", "U", &n, a, &lda, w, work.data(), &lwork, &info);
    return info;
}

std::vector<int> by_magnitude(const double *v, int n)
{
    std::vector<std::pair<int, double> > t;
    for (int i = 0; i < n; ++i)
        t.push_back(std::make_pair(i, v[i]));
    std::sort(t.begin(), t.end(), [](std::pair<int, double> const &a, std::pair<int, double> const &b)
              { return std::abs(a.second) > std::abs(b.second); });
    std::vector<int> out;
    for (auto &p : t) out.push_back(p.first);
    return out;
}

void small_mm(bool ta, bool tb, int m, int n, int k, const double *A, int lda, const double *B,
              int ldb, double *C, int ldc)
{
    for (int j = 0; j < n; ++j)
        for (int i = 0; i < m; ++i)
        {
            double s = 0.0;
            for (int l = 0; l < k; ++l)
                s += (ta ? A[l + (size_t)i * lda] : A[i + (size_t)l * lda]) *
                     (tb ? B[j + (size_t)l * ldb] : B[l + (size_t)j * ldb]);
            C[i + (size_t)j * ldc] = s;
        }
}

// grow a column-major block to `cols` columns, keeping its contents
void grow(vec &a, long long ld, int cols) { a.resize((size_t)ld * cols); }

// square host matrix with capacity: entries keep their (i, j) on resize
struct Square
{
    vec a; int ld = 0;
    void reserve(int n)
    {
        if (n <= ld) return;
        vec b((size_t)n * n, 0.0);
        for (int j = 0; j < ld; ++j)
            for (int i = 0; i < ld; ++i)
                b[i + (size_t)j * n] = a[i + (size_t)j * ld];
        a.swap(b); ld = n;
    }
    double &operator()(int i, int j) { return a[i + (size_t)j * ld]; }
};

// Generalised projected problem A X E + E X A' + C = 0, E = V'MV symmetric
// positive definite (MATLAB lyap(VAV, VBV, [], VMV), matlab/RAILSsolver.m:381,
// i.e. SLICOT sg03ad in matlab/mex/lyap.c:55-84).  Restated through the
// Cholesky factor E = L L':  (L^-1 A L^-T) Y + Y (L^-1 A L^-T)' + L^-1 C L^-T = 0,
// X = L^-T Y L^-1 -- deliberately a different reduction from the product's
// (which forms E^-1 A by LU), so that agreement is a real cross-check.
int generalized_dense_solve(int k, const double *A, int lda, const double *E, int lde,
                            const double *C, int ldc, double *X, int ldx)
{
    if (k == 0) return 0;
    vec L((size_t)k * k), At((size_t)k * k), Ct((size_t)k * k), Y((size_t)k * k);
    for (int j = 0; j < k; ++j)
        for (int i = 0; i < k; ++i)
        {
            // symmetrised: V'MV is symmetric up to rounding
            L[i + (size_t)j * k] = 0.5 * (E[i + (size_t)j * lde] + E[j + (size_t)i * lde]);
            At[i + (size_t)j * k] = A[i + (size_t)j * lda];
            Ct[i + (size_t)j * k] = C[i + (size_t)j * ldc];
        }
    int info = 0;
    dpotrf_("L", &k, L.data(), &k, &info);
    if (info) return 1000 + info;
    const double one = 1.0;
    // At <- L^-1 At L^-T, Ct <- L^-1 Ct L^-T
    dtrsm_("L", "L", "N", "N", &k, &k, &one, L.data(), &k, At.data(), &k);
    dtrsm_("R", "L", "T", "N", &k, &k, &one, L.data(), &k, At.data(), &k);
    dtrsm_("L", "L", "N", "N", &k, &k, &one, L.data(), &k, Ct.data(), &k);
    dtrsm_("R", "L", "T", "N", &k, &k, &one, L.data(), &k, Ct.data(), &k);
    info = oracle_dense_solve(k, At.data(), k, Ct.data(), k, Y.data(), k);
    // X = L^-T Y L^-1
    dtrsm_("L", "L", "T", "N", &k, &k, &one, L.data(), &k, Y.data(), &k);
    dtrsm_("R", "L", "N", "N", &k, &k, &one, L.data(), &k, Y.data(), &k);
    for (int j = 0; j < k; ++j)
        for (int i = 0; i < k; ++i)
            X[i + (size_t)j * ldx] = Y[i + (size_t)j * k];
    return info;
}

int solve_impl(
    int m, const int *rowptr, const int *colidx, const double *vals,
    const int *m_rowptr, const int *m_colidx, const double *m_vals,
    int n_b, const double *B_in, unsigned seed,
    int nparams, const char *const *keys, const double *pvals,
    int hist_cap, double *hist_abs, double *hist_rel, int *hist_k, double *hist_sec, int *hist_len,
    int *k_io, int v_cap_cols, double *V_io, double *T_out,
    double *solve_seconds, double *dense_solve_seconds)
{
    Params P;
    for (int i = 0; i < nparams; ++i) P.kv[keys[i]] = pvals[i];
    const int max_iter = (int)P.get("Maximum iterations", 1000);
    const double tol = P.get("Tolerance", 1e-3);
    const int expand_size = (int)P.get("Expand size", 3);
    const int L = (int)P.get("Lanczos iterations", 10);
    const int restart_size = (int)P.get("Restart size", -1);
    const int reduced_size = (int)P.get("Reduced size", -1);
    const int restart_iterations = (int)P.get("Restart iterations", 20);
    const double restart_tol = P.get("Restart tolerance", tol * 1e-3);
    const bool minimize = P.get("Minimize solution space", 1.0) != 0.0;
    const bool warm = P.get("Restart from solution", 0.0) != 0.0;
    // checker-only key (bench.py's bounded CPU samples): stop with code -2
    // after the first iteration that ends later than this many seconds
    const double time_limit = P.get("Oracle time limit", 0.0);
    // extension following matlab/RAILSsolver.m:368-392 (the C++ reference never
    // reads M): MV = M V, VMV = V'MV, generalised projected problem,
    // residual operator AV T MV' + MV T AV' + B B'.  PARITY UNPINNED: no
    // reference test fixes values for this branch.
    const bool use_mass = P.get("Use mass matrix", 0.0) != 0.0;
    if (L <= expand_size)
        return 101;
    if (use_mass && !m_rowptr)
        return 102;

    std::srand(seed);
    // a caller that made B itself with `srand(seed); B.random()` and passes it in
    // asks for that std::rand() value to be skipped, so that V and the Lanczos
    // start vectors get the values they get in a reference run
    for (int i = 0; i < (int)P.get("Skip random draws", 0.0); ++i) (void)std::rand();
    const long long ld = m;
    vec B((size_t)ld * n_b);
    if (B_in) std::memcpy(B.data(), B_in, B.size() * sizeof(double));
    else oracle_random(m, n_b, (unsigned)std::rand(), 0, m, B.data(), ld);

    int kV = warm && k_io ? *k_io : 1;
    int cap = std::max(kV, std::min(restart_size > 0 ? restart_size : 100, m));
    vec V((size_t)ld * cap), AV((size_t)ld * cap), BV((size_t)n_b * cap);
    if (!warm)
    {
        kV = 1;
        oracle_random(m, 1, (unsigned)std::rand(), 0, m, V.data(), ld);
        oracle_orthogonalize(m, 1, V.data(), ld, 0);
    }
    else
        std::memcpy(V.data(), V_io, (size_t)ld * kV * sizeof(double));
    int orth = warm ? 0 : 1;           // leading orthonormal columns of V
    if (warm) orth = kV;               // a previous solution basis is orthonormal

    int w0 = 0, we = kV;               // W = V(:, w0 : w0+we)
    int kAV = 0;
    Square VAV, VBV, VMV;
    VAV.reserve(cap); VBV.reserve(cap);
    vec MV;
    if (use_mass)
    {
        MV.resize((size_t)ld * cap);
        VMV.reserve(cap);
    }
    vec T;
    int kT = 0;

    double dense_sec = 0.0;
    auto t_begin = std::chrono::steady_clock::now();
    auto finish = [&](int code) {
        if (solve_seconds)
            *solve_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_begin).count();
        if (dense_solve_seconds) *dense_solve_seconds = dense_sec;
        if (k_io) *k_io = kV;
        if (V_io && T_out && kV <= v_cap_cols)
        {
            std::memcpy(V_io, V.data(), (size_t)ld * kV * sizeof(double));
            for (int j = 0; j < kV; ++j)
                for (int i = 0; i < kV; ++i)
                    T_out[i + (size_t)j * v_cap_cols] = T[i + (size_t)j * kT];
        }
        return code;
    };

    // r0 = |B|_2
    double r0;
    {
        vec g((size_t)n_b * n_b), w(n_b);
        oracle_dot(m, n_b, B.data(), ld, n_b, B.data(), ld, g.data());
        if (n_b == 1) r0 = std::sqrt(std::abs(g[0]));
        else
        {
            sym_eig(n_b, g.data(), n_b, w.data());
            r0 = 0.0;
            for (int i = 0; i < n_b; ++i) r0 = std::max(r0, std::sqrt(std::abs(w[i])));
        }
    }

    bool converged_before = false;
    int last_restart = 0, nhist = 0;
    vec Q((size_t)ld * (L + 1)), H((size_t)(L + 1) * (L + 1));

    for (int iter = 0; iter < max_iter; ++iter)
    {
        const int cols_before = kV;
        if (we)
        {
            // AW, BW; VAV = [[VAV; W'AV], V'AW]; VBV likewise (:141-207)
            const double *W = V.data() + (size_t)w0 * ld;
            double *AW = AV.data() + (size_t)kAV * ld;        // lands where push_back puts it
            oracle_spmm(m, rowptr, colidx, vals, we, W, ld, AW, ld);
            vec BW((size_t)n_b * we);
            oracle_dot(m, n_b, B.data(), ld, we, W, ld, BW.data());
            const int s = kAV + we;
            if (kAV > 0)
            {
                vec WAV((size_t)we * kAV), WBV((size_t)we * kAV);
                oracle_dot(m, we, W, ld, kAV, AV.data(), ld, WAV.data());
                small_mm(true, false, we, kAV, n_b, BW.data(), n_b, BV.data(), n_b, WBV.data(), we);
                for (int i = 0; i < we; ++i)
                    for (int j = 0; j < kAV; ++j)
                    {
                        VAV(i + kAV, j) = WAV[i + (size_t)j * we];
                        VBV(i + kAV, j) = WBV[i + (size_t)j * we];
                        VBV(j, i + kAV) = WBV[i + (size_t)j * we];
                    }
            }
            vec VAW((size_t)kV * we);
            oracle_dot(m, kV, V.data(), ld, we, AW, ld, VAW.data());
            for (int i = 0; i < kV; ++i)
                for (int j = 0; j < we; ++j)
                    VAV(i, j + kAV) = VAW[i + (size_t)j * kV];
            vec WBW((size_t)we * we);
            small_mm(true, false, we, we, n_b, BW.data(), n_b, BW.data(), n_b, WBW.data(), we);
            for (int i = 0; i < we; ++i)
                for (int j = 0; j < we; ++j)
                    VBV(i + kAV, j + kAV) = WBW[i + (size_t)j * we];
            std::memcpy(BV.data() + (size_t)kAV * n_b, BW.data(), BW.size() * sizeof(double));
            if (use_mass)
            {
                double *MW = MV.data() + (size_t)kAV * ld;
                oracle_spmm(m, m_rowptr, m_colidx, m_vals, we, W, ld, MW, ld);
                if (kAV > 0)
                {
                    vec WMV((size_t)we * kAV);
                    oracle_dot(m, we, W, ld, kAV, MV.data(), ld, WMV.data());
                    for (int i = 0; i < we; ++i)
                        for (int j = 0; j < kAV; ++j)
                            VMV(i + kAV, j) = WMV[i + (size_t)j * we];
                }
                vec VMW((size_t)kV * we);
                oracle_dot(m, kV, V.data(), ld, we, MW, ld, VMW.data());
                for (int i = 0; i < kV; ++i)
                    for (int j = 0; j < we; ++j)
                        VMV(i, j + kAV) = VMW[i + (size_t)j * kV];
            }
            kAV = s;
        }

        // T = projected solution
        {
            auto t0 = std::chrono::steady_clock::now();
            kT = kAV;
            T.assign((size_t)kT * kT, 0.0);
            int info = use_mass
                ? generalized_dense_solve(kT, VAV.a.data(), VAV.ld, VMV.a.data(), VMV.ld,
                                          VBV.a.data(), VBV.ld, T.data(), kT)
                : oracle_dense_solve(kT, VAV.a.data(), VAV.ld, VBV.a.data(), VBV.ld, T.data(), kT);
            if (info != 0 && info != kT + 1)
                std::cerr << "oracle: dense solve returned info = " << info << std::endl;
            dense_sec += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        }

        // residual Ritz pairs
        int steps = oracle_resid_lanczos(m, kV, use_mass ? MV.data() : V.data(), ld, AV.data(), ld,
                                         T.data(), kT, n_b,
                                         B.data(), ld, L, (unsigned)std::rand(), Q.data(), ld,
                                         H.data(), L + 1);
        vec Hs((size_t)steps * steps), theta(steps);
        for (int j = 0; j < steps; ++j)
            for (int i = 0; i < steps; ++i)
                Hs[i + (size_t)j * steps] = H[i + (size_t)j * (L + 1)];
        sym_eig(steps, Hs.data(), steps, theta.data());       // Hs now holds the Ritz coefficients
        double res = 0.0;
        for (int i = 0; i < steps; ++i) res = std::max(res, std::abs(theta[i]));

        double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_begin).count();
        if (nhist < hist_cap)
        {
            if (hist_abs) hist_abs[nhist] = res;
            if (hist_rel) hist_rel[nhist] = std::abs(res) / r0 / r0;
            if (hist_k) hist_k[nhist] = kV;
            if (hist_sec) hist_sec[nhist] = sec;
        }
        nhist++;
        if (hist_len) *hist_len = nhist;

        if (time_limit > 0.0 && sec > time_limit)
            return finish(-2);

        const bool converged = std::abs(res) < tol * r0 * r0;
        if (converged || iter + 1 >= max_iter || kV >= m)
        {
            if (converged && minimize && !converged_before)
                converged_before = true;
            else
                return finish(converged ? 0 : -1);
        }

        if ((restart_size > 0 && kV >= restart_size) ||
            (restart_iterations > 0 && iter - last_restart >= restart_iterations) || converged)
        {
            // restart basis: dominant eigenvectors of T (:449-482)
            vec E(T), lam(kT);
            sym_eig(kT, E.data(), kT, lam.data());
            int num = std::min(reduced_size, kV);
            num = num > 0 ? num : kT;
            std::vector<int> idx = by_magnitude(lam.data(), kT);
            vec X((size_t)kT * num);
            int kept = 0;
            for (int i = 0; i < num; ++i)
                if (std::abs(lam[idx[i]]) > restart_tol)
                {
                    for (int j = 0; j < kT; ++j)
                        X[j + (size_t)i * kT] = E[j + (size_t)idx[i] * kT];
                    kept++;
                }
            // V <- V X, AV <- AV X, BV <- BV X through a temporary (:265,290,297)
            vec tmp((size_t)ld * kept);
            oracle_gemm(m, kV, V.data(), ld, X.data(), kT, kept, 1.0, 0.0, tmp.data(), ld);
            std::memcpy(V.data(), tmp.data(), tmp.size() * sizeof(double));
            oracle_gemm(m, kAV, AV.data(), ld, X.data(), kT, kept, 1.0, 0.0, tmp.data(), ld);
            std::memcpy(AV.data(), tmp.data(), tmp.size() * sizeof(double));
            vec tb((size_t)n_b * kept);
            small_mm(false, false, n_b, kept, kAV, BV.data(), n_b, X.data(), kT, tb.data(), n_b);
            std::memcpy(BV.data(), tb.data(), tb.size() * sizeof(double));
            // VAV <- X' (VAV X), VBV <- X' (VBV X)
            vec t1((size_t)kT * kept), t2((size_t)kept * kept);
            small_mm(false, false, kT, kept, kT, VAV.a.data(), VAV.ld, X.data(), kT, t1.data(), kT);
            small_mm(true, false, kept, kept, kT, X.data(), kT, t1.data(), kT, t2.data(), kept);
            for (int j = 0; j < kept; ++j)
                for (int i = 0; i < kept; ++i) VAV(i, j) = t2[i + (size_t)j * kept];
            small_mm(false, false, kT, kept, kT, VBV.a.data(), VBV.ld, X.data(), kT, t1.data(), kT);
            small_mm(true, false, kept, kept, kT, X.data(), kT, t1.data(), kT, t2.data(), kept);
            for (int j = 0; j < kept; ++j)
                for (int i = 0; i < kept; ++i) VBV(i, j) = t2[i + (size_t)j * kept];
            if (use_mass)
            {
                oracle_gemm(m, kAV, MV.data(), ld, X.data(), kT, kept, 1.0, 0.0, tmp.data(), ld);
                std::memcpy(MV.data(), tmp.data(), tmp.size() * sizeof(double));
                small_mm(false, false, kT, kept, kT, VMV.a.data(), VMV.ld, X.data(), kT, t1.data(), kT);
                small_mm(true, false, kept, kept, kT, X.data(), kT, t1.data(), kT, t2.data(), kept);
                for (int j = 0; j < kept; ++j)
                    for (int i = 0; i < kept; ++i) VMV(i, j) = t2[i + (size_t)j * kept];
            }
            kV = kAV = kept;
            orth = std::min(orth, kept);
            we = 0;
            last_restart = iter;
            continue;
        }

        const int grow_by = std::min(std::min(expand_size, steps), (restart_size > 0 ? restart_size : m) - kV);
        if (kV + grow_by > cap)
        {
            cap += 100;
            grow(V, ld, cap); grow(AV, ld, cap); grow(BV, n_b, cap);
            VAV.reserve(cap); VBV.reserve(cap);
            if (use_mass) { grow(MV, ld, cap); VMV.reserve(cap); }
        }
        // new directions: Ritz vectors Q * y of the largest |theta|
        std::vector<int> idx = by_magnitude(theta.data(), steps);
        for (int i = 0; i < grow_by; ++i)
            oracle_gemm(m, steps, Q.data(), ld, Hs.data() + (size_t)idx[i] * steps, steps, 1, 1.0, 0.0,
                        V.data() + (size_t)(kV + i) * ld, ld);
        kV += grow_by;
        oracle_orthogonalize(m, kV, V.data(), ld, orth);
        orth = kV;
        w0 = cols_before; we = grow_by;
    }
    return finish(1);
}

}

extern "C" int oracle_solve(
    int m, const int *rowptr, const int *colidx, const double *vals,
    int n_b, const double *B_in, unsigned seed,
    int nparams, const char *const *keys, const double *pvals,
    int hist_cap, double *hist_abs, double *hist_rel, int *hist_k, double *hist_sec, int *hist_len,
    int *k_io, int v_cap_cols, double *V_io, double *T_out,
    double *solve_seconds, double *dense_solve_seconds)
{
    return solve_impl(m, rowptr, colidx, vals, nullptr, nullptr, nullptr, n_b, B_in, seed, nparams, keys,
                      pvals, hist_cap, hist_abs, hist_rel, hist_k, hist_sec, hist_len, k_io, v_cap_cols,
                      V_io, T_out, solve_seconds, dense_solve_seconds);
}

// same, with the CSR arrays of a mass matrix M (params "Use mass matrix" = 1)
extern "C" int oracle_solve_m(
    int m, const int *rowptr, const int *colidx, const double *vals,
    const int *m_rowptr, const int *m_colidx, const double *m_vals,
    int n_b, const double *B_in, unsigned seed,
    int nparams, const char *const *keys, const double *pvals,
    int hist_cap, double *hist_abs, double *hist_rel, int *hist_k, double *hist_sec, int *hist_len,
    int *k_io, int v_cap_cols, double *V_io, double *T_out,
    double *solve_seconds, double *dense_solve_seconds)
{
    return solve_impl(m, rowptr, colidx, vals, m_rowptr, m_colidx, m_vals, n_b, B_in, seed, nparams,
                      keys, pvals, hist_cap, hist_abs, hist_rel, hist_k, hist_sec, hist_len, k_io,
                      v_cap_cols, V_io, T_out, solve_seconds, dense_solve_seconds);
}
