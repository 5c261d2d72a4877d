// MatrixMarket file I/O for the command-line driver.
//
// The reference's executable reads its inputs with EpetraExt
// (/root/reference/src/main.cpp:66-68: MatrixMarketFileToCrsMatrix for A.mtx,
// B.mtx and M.mtx; :157-160 MatrixMarketFileToMultiVector for V.mtx / T.mtx)
// and writes the solution factors with MultiVectorToMatrixMarketFile
// (:126-129).  EpetraExt is not available here; these functions restate the
// two file formats it uses:
//
//   * sparse operators: `%%MatrixMarket matrix coordinate <field> <symmetry>`,
//     1-based `i j value` triples; entries given twice are summed (what
//     Epetra_CrsMatrix::InsertGlobalValues + FillComplete does), and the
//     symmetric / skew-symmetric storage schemes are expanded;
//   * dense blocks (V, T, eigenvectors): `%%MatrixMarket matrix array real
//     general`, a `rows cols` line and then one value per line in column-major
//     order, written with 17 significant digits so that a write/read round
//     trip is bit-exact.
//
// Host-only code: nothing here touches the GPU.
#ifndef RAILS_B200_MATRIXMARKET_HPP
#define RAILS_B200_MATRIXMARKET_HPP

#include <string>
#include <vector>

namespace RAILS
{
namespace io
{

// rows [row_begin, row_begin + rows) of a global_rows x cols matrix in CSR,
// GLOBAL column indices sorted inside each row, int32 indices like the
// reference (`int` everywhere)
struct CsrHost
{
    int rows = 0;          // local rows
    int global_rows = 0;
    int cols = 0;
    int row_begin = 0;
    std::vector<int> rowptr;
    std::vector<int> colidx;
    std::vector<double> vals;
};

// All functions return 0 on success; on failure they return non-zero and put
// a one-line description in `err` (no exceptions, like the reference).

// Sparse matrix from a coordinate file.  Only the rows
// [row_begin, row_end) are kept (row_end < 0: all rows), so that each rank of
// a row-partitioned run reads the same file and keeps its own block.
int read_crs(std::string const &path, CsrHost &out, std::string &err,
             long long row_begin = 0, long long row_end = -1);

// Size line of a file (array or coordinate), without reading the entries
int read_size(std::string const &path, long long &rows, long long &cols, std::string &err);

// Dense block, column-major with leading dimension = rows.  Accepts the array
// format and, for convenience, a coordinate file (missing entries are zero).
int read_dense(std::string const &path, int &rows, int &cols, std::vector<double> &data,
               std::string &err);

// Dense block in array format (the layout of
// EpetraExt::MultiVectorToMatrixMarketFile); `ld` is the leading dimension of
// `data`.  `comment` (may be empty) is written as a `%` line after the banner.
int write_dense(std::string const &path, int rows, int cols, const double *data, long long ld,
                std::string const &comment, std::string &err);

// Sparse matrix in coordinate format (general, real)
int write_crs(std::string const &path, CsrHost const &A, std::string &err);

// true when the matrix is square with a unit diagonal and nothing else
bool is_identity(CsrHost const &A);
// true when every stored entry sits on the diagonal
bool is_diagonal(CsrHost const &A);
// number of rows whose diagonal entry is absent or exactly zero
long long zero_diagonal_rows(CsrHost const &A);

}
}

#endif
