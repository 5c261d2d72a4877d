// RAILS::b200::Solver -- the restarted projection loop for
//     A V T V' + V T V' A' + B B' = 0,      X ~ V T V'
// over the same three duck-typed concepts as the reference's
// RAILS::Solver<Matrix, MultiVector, DenseMatrix>
// (/root/reference/src/LyapunovSolverDecl.hpp:9-51).
//
// Why it exists: the reference loop is header-only C++ that lives under
// /root/reference and does not travel; this is the loop the product ships so
// that the CUDA wrappers can run where the reference tree is absent.  It is a
// from-scratch restatement of the ALGORITHM of src/LyapunovSolver.hpp:100-482
// -- same public members, same parameter names and defaults (:27-36,76-87),
// same return codes (0 converged / -1 stopped / 1 fell through, :239-240,345),
// same progress lines on std::cout (:219-221,232-268) -- and it issues the
// backend operations in the same order, so on any given backend it yields the
// same bits as the reference loop (tests/test_solver_loop.py checks exactly
// that against oracle/_ref with the reference's own StlWrapper).
//
// Additions that do not change the arithmetic: a per-iteration record
// (history()), an optional observer called once per outer iteration (bench.py
// uses it to bracket timed iterations with CUDA events) and wall-clock section
// totals (sections()) with the reference's timer names.
#ifndef RAILS_B200_RAILSSOLVER_HPP
#define RAILS_B200_RAILSSOLVER_HPP

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <functional>
#include <iostream>
#include <locale>
#include <map>
#include <string>
#include <utility>
#include <vector>

extern "C" {
void sb03md_(char *dico, char *job, char *fact, char *trans, int *n, double *A, int *lda,
             double *U, int *ldu, double *X, int *ldx, double *scale, double *sep,
             double *ferr, double *alphar, double *alphai, int *iwork, double *dwork,
             int *ldwork, int *info);
void dsyev_(char *jobz, char *uplo, int *n, double *a, int *lda, double *w, double *work,
            int *ldwork, int *info);
void dgesv_(int *n, int *nrhs, double *a, int *lda, int *ipiv, double *b, int *ldb, int *info);
}

namespace RAILS
{
namespace b200
{

// B may be an Operator or a MultiVector (src/MatrixOrMultiVectorWrapper.hpp:7-61)
template<class Matrix, class MultiVector>
class InputB
{
    bool is_matrix_;
    Matrix matrix_;
    MultiVector vector_;

public:
    InputB(Matrix const &b) : is_matrix_(true), matrix_(b) {}
    InputB(MultiVector const &b) : is_matrix_(false), vector_(b) {}

    double norm() const { return is_matrix_ ? matrix_.norm() : vector_.norm(); }
    MultiVector apply(MultiVector const &x) const
    { return is_matrix_ ? matrix_ * x : vector_ * x; }
    MultiVector apply_transposed(MultiVector const &x) const
    { return is_matrix_ ? matrix_.transpose() * x : vector_.transpose() * x; }
    bool is_multivector() const { return !is_matrix_; }
    MultiVector const &multivector() const { return vector_; }
};

// Matrix and MultiVector are one type (StlWrapper, ScalarWrapper)
template<class Type>
class InputB<Type, Type>
{
    Type b_;

public:
    template<class Any>
    InputB(Any const &b) : b_(b) {}

    double norm() const { return b_.norm(); }
    Type apply(Type const &x) const { return b_ * x; }
    Type apply_transposed(Type const &x) const { return b_.transpose() * x; }
    bool is_multivector() const { return true; }
    Type const &multivector() const { return b_; }
};

// Customisation point: a backend may supply the whole Lanczos recurrence of
// resid_lanczos as fused kernels.  The primary template is disabled, so any
// backend (StlWrapper, ScalarWrapper, ...) runs the operation-by-operation
// sequence of the reference; CudaFusedLanczos.hpp specialises it for the CUDA
// wrappers.  run() fills H (steps x steps tridiagonal) and Q (m x steps) and
// returns the number of steps taken.
template<class MultiVector, class DenseMatrix>
struct FusedLanczos
{
    static const bool enabled = false;
    static bool active() { return false; }
    // whether a space of k columns with an n-column B fits the fused kernels
    static bool fits(int, int) { return true; }
    static int run(MultiVector const &, MultiVector const &, DenseMatrix const &,
                   MultiVector const &, DenseMatrix &, MultiVector &, int) { return -1; }
    // optional head start before T is known (Q holds the random start vector)
    static void begin(MultiVector const &, MultiVector const &, MultiVector const &, MultiVector &, int) {}
};

// Customisation point: several independent inner products X_i' Y_i returned
// with one host round trip.  The primary template just calls dot() in order.
template<class MultiVector, class DenseMatrix>
struct BatchedDots
{
    static void run(std::vector<std::pair<MultiVector const *, MultiVector const *> > const &pairs,
                    std::vector<DenseMatrix> &out)
    {
        out.clear();
        for (size_t i = 0; i < pairs.size(); ++i)
            out.push_back(pairs[i].first->dot(*pairs[i].second));
    }
};

// Customisation point: failed backend calls so far.  The wrappers follow the
// reference's style (print to std::cerr, return an unfilled result); a backend
// that can fail at run time (device out of memory) reports it here and the loop
// stops with return code 2 instead of iterating on garbage.
template<class MultiVector>
struct BackendHealth
{
    static long long errors() { return 0; }
    // Called where a timed section closes.  A backend whose calls are asynchronous
    // waits here for its queued work WHEN PROFILING IS ON (RAILS_PROFILE=1 for the CUDA
    // wrappers), so that the section table shows where device time goes rather than
    // where the host happened to block; a no-op otherwise.
    static void fence() {}
};

struct IterationRecord
{
    int iteration;        // 1-based
    int space_size;       // columns of V when the residual was estimated
    double residual;      // largest Ritz value of the residual (absolute)
    double relative;      // residual / |B|_2^2
    bool restarted;       // this iteration ended in a restart
    double seconds;       // wall clock since solve() began
};

template<class Matrix, class MultiVector, class DenseMatrix>
class Solver
{
public:
    typedef std::function<void(IterationRecord const &)> Observer;

    template<class MatrixOrMultiVector>
    Solver(Matrix const &A, MatrixOrMultiVector const &B, Matrix const &M)
        : A_(A), B_(B), M_(M),
          max_iter_(1000), tol_(1e-3), expand_size_(3), lanczos_iterations_(10),
          restart_size_(-1), reduced_size_(-1), restart_iterations_(20),
          restart_tolerance_(tol_ * 1e-3), minimize_solution_space_(true),
          restart_from_solution_(false), use_mass_(false), residual_method_(0),
          m_orthogonalization_(false), relative_restart_tolerance_(false), restart_upon_start_(false),
          orthogonalize_candidates_(false), lanczos_vectors_(-1), has_nullspace_(false),
          projection_method_(1.0), has_inverse_(false),
          verbose_(true)
    {}

    virtual ~Solver() {}

    template<class ParameterList>
    int set_parameters(ParameterList &params);

    // Solve A*V*T*V' + V*T*V'*A' + B*B' = 0
    int solve(MultiVector &V, DenseMatrix &T);

    // Solve A*X + X*A' + B = 0 on the projected (k x k) problem
    int dense_solve(DenseMatrix const &A, DenseMatrix const &B, DenseMatrix &X);

    // Ritz pairs of the residual R = AV*T*V' + V*T*AV' + B*B'
    int resid_lanczos(MultiVector const &AV, MultiVector const &V, DenseMatrix const &T,
                      DenseMatrix &H, MultiVector &eigenvectors, DenseMatrix &eigenvalues,
                      int max_iter);
    // same, continuing from a start vector that lanczos_start() already drew
    // (and, on a backend with fused kernels, already projected)
    int resid_lanczos(MultiVector const &AV, MultiVector const &V, DenseMatrix const &T,
                      DenseMatrix &H, MultiVector &eigenvectors, DenseMatrix &eigenvalues,
                      int max_iter, MultiVector *started);
    // Q = random start vector (:372-374), drawn BEFORE the projected solve so
    // that a backend can overlap the first projection with it; returns false
    // when the backend has no use for that
    bool lanczos_start(MultiVector const &AV, MultiVector const &V, MultiVector &Q, int max_iter);

    // Exact eigenpairs of the residual through a tall-skinny QR of its factor
    // ("Residual method" = 1; BASELINE.json north_star item (3)):
    //     R = F S F',  F = [V, AV, B] (m x (2k+n)),  S = [[0 T 0]; [T 0 0]; [0 0 I]]
    //     F = Q_F R_F  ->  eig(R_F S R_F') = Y Theta Y'  ->  eigenvectors Q_F Y
    // No random start vector and no Lanczos breakdown: the result is the same
    // to rounding on every backend and GPU count.
    // The factorisation is kept in `state` across outer iterations and only the
    // columns that are new since the last call are orthogonalised.
    struct TsqrState
    {
        MultiVector Q;            // orthonormal basis of span F, r columns
        DenseMatrix RV, RA, RB;   // r x k, r x k, r x n:  [V, AV, B] = Q [RV, RA, RB]
        DenseMatrix RM;           // r x k: MV = Q RM ("Residual method" = 2 with a mass matrix)
        int columns;              // columns of V / AV already factorised
        int capacity;             // column capacity of Q
        bool valid;
        TsqrState() : columns(0), capacity(0), valid(false) {}
    };
    int resid_tsqr(TsqrState &state, MultiVector const &AV, MultiVector const &V,
                   DenseMatrix const &T, MultiVector &eigenvectors, DenseMatrix &eigenvalues,
                   int want);
    void tsqr_append(TsqrState &state, std::vector<MultiVector const *> const &blocks,
                     std::vector<DenseMatrix *> const &Rs);
    // "Residual method" = 2: the whole iteration in the coordinates of that basis
    // (RailsCoefficientSpace.hpp)
    int solve_coefficient_space(MultiVector &V, DenseMatrix &T);
    void tsqr_shrink(TsqrState &state);

    // Restart basis from the dominant eigenvectors of T
    int compute_restart_vectors(DenseMatrix &X, DenseMatrix const &T, int num, double tol);

    // Generalised projected problem A X E' + E X A' + B = 0 ("Use mass matrix";
    // MATLAB lyap(VAV, VBV, [], VMV), matlab/RAILSsolver.m:381)
    int dense_solve_generalized(DenseMatrix const &A, DenseMatrix const &E, DenseMatrix const &B,
                                DenseMatrix &X);

    // ---- options of the MATLAB implementation (matlab/RAILSsolver.m), all off
    // by default so that the loop above stays the C++ reference's ---------------
    // opts.nullspace (:28-29,218-220): a space projected out of every vector
    // that enters V.  Orthonormalised here (Morth(opts.nullspace, [])).
    void set_nullspace(MultiVector const &N);
    // Morth (:538-617): orthonormalise the columns of W against `against` and
    // V -- Euclidean inner product, or the M inner product with
    // "M-orthogonalization" -- and append the survivors (norm >= 1e-8 after the
    // projection) to V.  Two classical Gram-Schmidt passes per column in place
    // of MATLAB's single modified pass: same result to rounding, and each pass
    // is one V'w + one V*c kernel instead of k dependent axpys.  Returns the
    // number of columns appended.
    int morth(MultiVector const &W, MultiVector const *against, MultiVector &V);
    // opts.Ainv / opts.invA (:19-24,155-160): an operator that applies A^-1,
    // possibly approximately.  Only read when "Projection method" > 1 (:7-17):
    //   1.x  the space is expanded with A^-1 r instead of the residual vectors r,
    //   2.x  with [r, A^-1 r] (the extended Krylov flavour);
    //   x.1  start from A^-1 V0, x.2 from B (and A^-1 B), x.3 from V0 alone
    // (:294-308,520-524).  For a singular mass matrix the reference pairs this
    // with the Schur-complement operator and its Sinv (matlab/RAILSschur.m:60-66,
    // matlab/test/test_MOC.m:45-49).
    void set_inverse(Matrix const &Ainv) { Ainv_ = Ainv; has_inverse_ = true; }

    // ---- additions ------------------------------------------------------
    void set_observer(Observer obs) { observer_ = obs; }
    void set_verbose(bool v) { verbose_ = v; }
    std::vector<IterationRecord> const &history() const { return history_; }
    // accumulated host wall-clock seconds per section (reference timer names)
    std::map<std::string, double> const &sections() const { return sections_; }

protected:
    Matrix A_;
    InputB<Matrix, MultiVector> B_;
    // The reference C++ loop stores M and never reads it.  With
    // "Use mass matrix" = true this loop follows the MATLAB implementation
    // instead (matlab/RAILSsolver.m:368-392): MV = M V, VMV = V'MV, the projected
    // problem is the generalised one and the residual operator is
    // AV T MV' + MV T AV' + B B'.
    Matrix M_;

    int max_iter_;
    double tol_;
    int expand_size_;
    int lanczos_iterations_;
    int restart_size_;
    int reduced_size_;
    int restart_iterations_;
    double restart_tolerance_;
    bool minimize_solution_space_;
    bool restart_from_solution_;
    bool use_mass_;
    int residual_method_;     // 0: the reference's Lanczos, 1: TSQR + small eigenproblem,
                              // 2: the whole iteration in the coordinates of the TSQR basis
    // MATLAB-only behaviour, each with the option it restates
    bool m_orthogonalization_;        // opts.ortho = 'M'        (:31-34,303-304,406,527-528)
    bool relative_restart_tolerance_; // d / max(d) > tol        (:476-477)
    bool restart_upon_start_;         // opts.restart_upon_start (:48-51,456)
    bool orthogonalize_candidates_;   // Morth(V2, .., V) before the selection (:403-409,517-531)
    int lanczos_vectors_;             // opts.lanczos_vectors: candidates kept (:72-77,309-311)
    bool has_nullspace_;
    MultiVector nullspace_;
    double projection_method_;        // opts.projection_method  (:7-17,294-308,520-524)
    bool has_inverse_;
    Matrix Ainv_;
    bool verbose_;

    Observer observer_;
    std::vector<IterationRecord> history_;
    std::map<std::string, double> sections_;

    struct Section
    {
        std::map<std::string, double> &acc;
        std::string name;
        std::chrono::steady_clock::time_point t0;
        Section(std::map<std::string, double> &a, const char *n)
            : acc(a), name(n), t0(std::chrono::steady_clock::now()) {}
        ~Section()
        {
            BackendHealth<MultiVector>::fence();
            acc[name] += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        }
    };

    void project(MultiVector const &V, MultiVector const &W, MultiVector &AV, MultiVector &BV,
                 MultiVector &MV, DenseMatrix &VAV, DenseMatrix &VBV, DenseMatrix &VMV, bool first,
                 int max_size);
    void compress(MultiVector &V, MultiVector &W, MultiVector &AV, MultiVector &BV, MultiVector &MV,
                  DenseMatrix &VAV, DenseMatrix &VBV, DenseMatrix &VMV, DenseMatrix const &T,
                  TsqrState *tsqr = nullptr);
};

// A key is looked up as given, then UPPER, lower and Title Case; the last
// spelling that is present wins (src/LyapunovSolver.hpp:40-70).
template<class ParameterList, class Type>
Type lookup_parameter(ParameterList &params, std::string const &name, Type def)
{
    std::locale loc;
    Type value = params.get(name, def);

    std::string upper(name), lower(name), title(name);
    for (size_t i = 0; i < name.size(); ++i)
    {
        upper[i] = std::toupper(name[i], loc);
        lower[i] = std::tolower(name[i], loc);
    }
    if (!title.empty())
        title[0] = std::toupper(title[0]);
    for (size_t i = 1; i < title.size(); ++i)
        if (!isalpha(title[i - 1]) && islower(title[i]))
            title[i] = std::toupper(title[i], loc);

    value = params.get(upper, value);
    value = params.get(lower, value);
    value = params.get(title, value);
    return value;
}

template<class Matrix, class MultiVector, class DenseMatrix>
template<class ParameterList>
int Solver<Matrix, MultiVector, DenseMatrix>::set_parameters(ParameterList &params)
{
    max_iter_ = lookup_parameter(params, "Maximum iterations", max_iter_);
    tol_ = lookup_parameter(params, "Tolerance", tol_);
    expand_size_ = lookup_parameter(params, "Expand size", expand_size_);
    lanczos_iterations_ = lookup_parameter(params, "Lanczos iterations", lanczos_iterations_);
    restart_size_ = lookup_parameter(params, "Restart size", restart_size_);
    reduced_size_ = lookup_parameter(params, "Reduced size", reduced_size_);
    restart_iterations_ = lookup_parameter(params, "Restart iterations", restart_iterations_);
    // recomputed from the NEW tolerance unless given (:83)
    restart_tolerance_ = lookup_parameter(params, "Restart tolerance", tol_ * 1e-3);
    minimize_solution_space_ = lookup_parameter(params, "Minimize solution space",
                                                minimize_solution_space_);
    restart_from_solution_ = lookup_parameter(params, "Restart from solution",
                                              restart_from_solution_);
    // extension (not a reference key): generalised equation with the stored M
    use_mass_ = lookup_parameter(params, "Use mass matrix", use_mass_);
    // extension: 0 = Lanczos as the reference, 1 = TSQR of [V, AV, B] + small eigenproblem,
    // 2 = projection, residual and orthogonalisation all in the coordinates of that basis
    residual_method_ = lookup_parameter(params, "Residual method", residual_method_);

    // extensions: options of matlab/RAILSsolver.m (see the member declarations)
    m_orthogonalization_ = lookup_parameter(params, "M-orthogonalization", m_orthogonalization_);
    relative_restart_tolerance_ = lookup_parameter(params, "Relative restart tolerance",
                                                   relative_restart_tolerance_);
    restart_upon_start_ = lookup_parameter(params, "Restart upon start", restart_upon_start_);
    orthogonalize_candidates_ = lookup_parameter(params, "Orthogonalize candidates",
                                                 orthogonalize_candidates_);
    lanczos_vectors_ = lookup_parameter(params, "Lanczos vectors", lanczos_vectors_);
    projection_method_ = lookup_parameter(params, "Projection method", projection_method_);
    if (projection_method_ < 1.0 || projection_method_ >= 3.0)
    {
        std::cerr << "Projection method " << projection_method_ << " does not exist "
                  << "(1, 1.1, 1.2, 1.3, 2.1, 2.2, 2.3)" << std::endl;
        return 1;
    }
    if (m_orthogonalization_)
        use_mass_ = true;          // V'MV = I: the residual still needs MV (:391)

    if (lanczos_iterations_ <= expand_size_)
    {
        std::cerr << "Amount of Lanczos iterations is smaller than "
                  << "the amount of vectors that are used to expand "
                  << "the space in every iteration" << std::endl;
        return 1;
    }
    return 0;
}

}
}

#include "RailsSolverImpl.hpp"
#include "RailsCoefficientSpace.hpp"

#endif
