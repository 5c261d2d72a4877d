// C entry points of librails_host.so for the pieces around the solver loop
// (SURVEY.md §8(f) N1, N2): MatrixMarket files, parameter files and the
// eigenpairs / trace of a low-rank solution.  Plain pointers and sizes, int
// return codes, the message of the last failure kept for
// rails_host_last_message().  The file and parameter functions are host-only;
// rails_host_solution_eigs needs an initialised GPU (no CPU stand-in).
#include "CudaMultiVectorWrapper.hpp"
#include "HostDenseMatrix.hpp"
#include "MatrixMarket.hpp"
#include "ParameterFile.hpp"
#include "SolutionTools.hpp"

#include <algorithm>
#include <cstring>
#include <iostream>
#include <string>

using namespace RAILS;

namespace
{
std::string g_message;
io::CsrHost g_crs;            // result of the last rails_host_mm_read_crs
std::vector<double> g_dense;  // result of the last rails_host_mm_read_dense
}

extern "C" {

const char *rails_host_last_message(void) { return g_message.c_str(); }

// ---- MatrixMarket ---------------------------------------------------------
// Two steps so that the caller can size its arrays: read (dims = rows,
// global_rows, cols, nnz of the kept row block), then fetch.
int rails_host_mm_read_crs(const char *path, long long row_begin, long long row_end, long long dims[4])
{
    g_crs = io::CsrHost();
    if (io::read_crs(path, g_crs, g_message, row_begin, row_end))
        return 1;
    dims[0] = g_crs.rows;
    dims[1] = g_crs.global_rows;
    dims[2] = g_crs.cols;
    dims[3] = (long long)g_crs.vals.size();
    return 0;
}

int rails_host_mm_fetch_crs(int *rowptr, int *colidx, double *vals)
{
    std::copy(g_crs.rowptr.begin(), g_crs.rowptr.end(), rowptr);
    std::copy(g_crs.colidx.begin(), g_crs.colidx.end(), colidx);
    std::copy(g_crs.vals.begin(), g_crs.vals.end(), vals);
    return 0;
}

int rails_host_mm_read_size(const char *path, long long *rows, long long *cols)
{
    return io::read_size(path, *rows, *cols, g_message);
}

int rails_host_mm_read_dense(const char *path, int *rows, int *cols)
{
    g_dense.clear();
    return io::read_dense(path, *rows, *cols, g_dense, g_message);
}

int rails_host_mm_fetch_dense(double *data)
{
    std::copy(g_dense.begin(), g_dense.end(), data);
    return 0;
}

int rails_host_mm_write_dense(const char *path, int rows, int cols, const double *data, long long ld,
                              const char *comment)
{
    return io::write_dense(path, rows, cols, data, ld, comment ? comment : "", g_message);
}

int rails_host_mm_write_crs(const char *path, int rows, int global_rows, int cols, int row_begin,
                            const int *rowptr, const int *colidx, const double *vals)
{
    io::CsrHost A;
    A.rows = rows;
    A.global_rows = global_rows;
    A.cols = cols;
    A.row_begin = row_begin;
    A.rowptr.assign(rowptr, rowptr + rows + 1);
    A.colidx.assign(colidx, colidx + rowptr[rows]);
    A.vals.assign(vals, vals + rowptr[rows]);
    return io::write_crs(path, A, g_message);
}

// ---- parameter files -------------------------------------------------------
// value of `key` in sublist `sublist` ("" or NULL: top level) of the file, as
// a double (booleans as 0/1); *found = 0 and *value untouched when absent
int rails_host_params_get(const char *path, const char *sublist, const char *key, double *value,
                          int *found)
{
    ParameterTree tree;
    if (ParameterTree::from_file(path, tree, g_message))
        return 1;
    ParameterTree const *list = &tree;
    if (sublist && *sublist)
    {
        list = tree.find_sublist(sublist);
        if (!list)
        {
            *found = 0;
            return 0;
        }
    }
    *found = list->isParameter(key) ? 1 : 0;
    if (*found)
        *value = list->get(key, *value);
    return 0;
}

// the file re-emitted as Teuchos-style XML (NUL-terminated, truncated to cap)
int rails_host_params_to_xml(const char *path, char *out, int cap)
{
    ParameterTree tree;
    if (ParameterTree::from_file(path, tree, g_message))
        return 1;
    std::string xml = tree.to_xml();
    if (cap > 0)
    {
        size_t n = std::min((size_t)cap - 1, xml.size());
        std::memcpy(out, xml.data(), n);
        out[n] = '\0';
    }
    return 0;
}

// ---- eigenpairs and trace of X = V T V' ------------------------------------
// V_host: this rank's rows of V (rows x k, leading dimension ldv); T_host k x k.
// On return *found pairs: eigenvalues[0..found), eigenvectors (rows x found,
// leading dimension ldq) and trace(X).  Mirrors Epetra_OperatorWrapper::eigs +
// SchurOperator::Trace as used by src/main.cpp:140-170.
int rails_host_solution_eigs(int rows, int global_rows, int k, const double *V_host, long long ldv,
                             const double *T_host, int ldt, int num, double tol, int *found,
                             double *eigenvalues, double *eigenvectors, long long ldq, double *trace)
{
    if (!rails_cuda_is_initialized())
    {
        g_message = "rails_cuda_init has not been called";
        return 200;
    }
    CudaMultiVectorWrapper V(global_rows, k);
    if (k > 0 && rails_mv_upload(V.device_handle(), 0, k, V_host, ldv))
    {
        g_message = rails_cuda_last_error();
        return 200;
    }
    (void)rows;
    HostDenseMatrix T(k, k);
    for (int j = 0; j < k; ++j)
        for (int i = 0; i < k; ++i)
            T(i, j) = T_host[i + (size_t)j * ldt];
    CudaMultiVectorWrapper Q;
    HostDenseMatrix lambda(1, 1);
    int rc = b200::solution_eigs(V, T, Q, lambda, num, tol);
    if (rc)
    {
        g_message = "solution_eigs failed";
        return rc;
    }
    *found = lambda.M();
    for (int i = 0; i < lambda.M(); ++i)
        eigenvalues[i] = lambda(i, 0);
    if (lambda.M() > 0 && eigenvectors && Q.to_host(eigenvectors, ldq))
    {
        g_message = rails_cuda_last_error();
        return 200;
    }
    if (trace)
        *trace = b200::solution_trace(V, T);
    return 0;
}

}
