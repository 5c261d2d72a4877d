// Sparse LU factorisation on the host, the set-up step of the Schur-complement
// operator for a singular mass matrix.
//
// The reference factorises A11 once with Amesos_Klu (third-party, absent here:
// /root/reference/src/SchurOperator.cpp:33,177-186) and MATLAB's RAILSschur
// with lu() (matlab/RAILSschur.m:31-33); every operator apply then solves with
// the factors.  Here the factorisation is a from-scratch left-looking
// (Gilbert-Peierls) LU with threshold partial pivoting that prefers the
// diagonal, after a fill-reducing ordering of A + A': nested dissection by
// breadth-first level structures or minimum degree, whichever fills less (KLU
// uses AMD after a block-triangular pre-pass).  It runs once per operator on the host, like
// the reference's; the solves -- the per-iteration work -- run on the GPU from
// the factors uploaded by rails_lu_create (csrc/sptrsv.cu).
//
//     P A Q = L U      L unit lower triangular (strict part stored),
//                      U upper triangular with its diagonal,
//     both CSC in pivot numbering; pinv[i] = pivot position of original row i,
//     q[j] = original column eliminated at step j.
#ifndef RAILS_B200_SPARSELU_HPP
#define RAILS_B200_SPARSELU_HPP

#include <vector>

namespace RAILS
{

struct SparseLU
{
    int n = 0;
    std::vector<int> Lp, Li;
    std::vector<double> Lx;
    std::vector<int> Up, Ui;
    std::vector<double> Ux;
    std::vector<int> pinv, q;
    // default ordering: nested dissection; up to this many rows minimum degree is
    // computed as well and the ordering with the smaller symbolic fill is used
    int compare_orderings_up_to = 8000;
    int off_diagonal_pivots = 0;      // steps where the diagonal failed the threshold
    double min_abs_pivot = 0.0, max_abs_pivot = 0.0;

    long long nnz() const { return (long long)Li.size() + (long long)Ui.size(); }

    // Factorises the n x n CSR matrix.  `order` (n entries, a permutation) fixes
    // the elimination order of the columns; nullptr = the default described above.
    // Returns 0, or j + 1 when no non-zero pivot exists at step j (singular).
    int factor(int n, const int *rowptr, const int *colidx, const double *vals,
               double pivot_threshold = 0.1, const int *order = nullptr);

    // x := A^-1 x (trans = false) or A'^-1 x (trans = true) for nrhs columns of
    // leading dimension ld.  Host arithmetic: used by the set-up checks and the
    // CPU tests of the factorisation, never by an operator apply.
    void solve(double *x, int nrhs, long long ld, bool trans = false) const;
};

// Minimum-degree ordering of the graph of A + A' (rows of more than
// max(16, 10 sqrt(n)) entries are left to the end, as AMD does with dense rows).
std::vector<int> minimum_degree_order(int n, const int *rowptr, const int *colidx);

// Nested dissection by breadth-first level structures, minimum degree on the
// pieces of at most `leaf` nodes.
std::vector<int> nested_dissection_order(int n, const int *rowptr, const int *colidx, int leaf = 200);

// Entries below the diagonal of the Cholesky factor of the pattern of A + A' under
// `order`: the fill to expect in L and in U.
long long symbolic_fill(int n, const int *rowptr, const int *colidx, const int *order);

}

#endif
