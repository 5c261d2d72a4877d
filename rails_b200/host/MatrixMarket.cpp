#include "MatrixMarket.hpp"

#include <exception>

#include <algorithm>
#include <cctype>
#include <cerrno>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <sstream>

namespace RAILS
{
namespace io
{

namespace
{

struct Banner
{
    bool coordinate = false;   // else array
    bool pattern = false;      // no value column
    bool complex_field = false;
    int symmetry = 0;          // 0 general, 1 symmetric, 2 skew-symmetric
};

std::string lower(std::string s)
{
    for (size_t i = 0; i < s.size(); ++i)
        s[i] = (char)std::tolower((unsigned char)s[i]);
    return s;
}

// whole file in memory, NUL-terminated so that strtod/strtoll stop at the end
int slurp(std::string const &path, std::vector<char> &buf, std::string &err)
{
    FILE *f = std::fopen(path.c_str(), "rb");
    if (!f)
    {
        err = "cannot open " + path + ": " + std::strerror(errno);
        return 1;
    }
    std::fseek(f, 0, SEEK_END);
    long long size = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    buf.resize((size_t)size + 1);
    size_t got = size > 0 ? std::fread(buf.data(), 1, (size_t)size, f) : 0;
    std::fclose(f);
    if ((long long)got != size)
    {
        err = "short read on " + path;
        return 1;
    }
    buf[(size_t)size] = '\0';
    return 0;
}

// parses the banner and skips the comment block; `p` is left on the size line
int parse_banner(std::string const &path, const char *&p, Banner &b, std::string &err)
{
    const char *eol = std::strchr(p, '\n');
    std::string first(p, eol ? (size_t)(eol - p) : std::strlen(p));
    std::istringstream is(first);
    std::string tag, object, format, field, symmetry;
    is >> tag >> object >> format >> field >> symmetry;
    if (lower(tag) != "%%matrixmarket" || lower(object) != "matrix")
    {
        err = path + ": not a MatrixMarket matrix file (banner: '" + first + "')";
        return 1;
    }
    format = lower(format);
    field = lower(field);
    symmetry = lower(symmetry);
    if (format == "coordinate")
        b.coordinate = true;
    else if (format != "array")
    {
        err = path + ": unknown format '" + format + "'";
        return 1;
    }
    if (field == "pattern")
        b.pattern = true;
    else if (field == "complex")
        b.complex_field = true;
    else if (field != "real" && field != "integer" && field != "double")
    {
        err = path + ": unknown field '" + field + "'";
        return 1;
    }
    if (b.complex_field)
    {
        err = path + ": complex matrices are not supported";
        return 1;
    }
    if (symmetry == "general" || symmetry.empty())
        b.symmetry = 0;
    else if (symmetry == "symmetric")
        b.symmetry = 1;
    else if (symmetry == "skew-symmetric")
        b.symmetry = 2;
    else
    {
        err = path + ": unsupported symmetry '" + symmetry + "'";
        return 1;
    }
    p = eol ? eol + 1 : p + std::strlen(p);
    // comment and blank lines
    for (;;)
    {
        const char *q = p;
        while (*q == ' ' || *q == '\t' || *q == '\r')
            ++q;
        if (*q == '%' || *q == '\n')
        {
            const char *e = std::strchr(q, '\n');
            if (!e)
            {
                p = q + std::strlen(q);
                break;
            }
            p = e + 1;
            continue;
        }
        break;
    }
    return 0;
}

bool next_int(const char *&p, long long &v)
{
    char *e = nullptr;
    errno = 0;
    v = std::strtoll(p, &e, 10);
    if (e == p)
        return false;
    p = e;
    return true;
}

bool next_double(const char *&p, double &v)
{
    char *e = nullptr;
    v = std::strtod(p, &e);
    if (e == p)
        return false;
    p = e;
    return true;
}

struct Triple
{
    int row;       // local
    int col;
    double val;
};

}

static int read_size_unguarded(std::string const &path, long long &rows, long long &cols, std::string &err)
{
    FILE *f = std::fopen(path.c_str(), "rb");
    if (!f)
    {
        err = "cannot open " + path + ": " + std::strerror(errno);
        return 1;
    }
    // the banner and the comment block are short; 1 MiB is far more than any
    // writer produces before the size line
    std::vector<char> buf((1 << 20) + 1);
    size_t got = std::fread(buf.data(), 1, buf.size() - 1, f);
    std::fclose(f);
    buf[got] = '\0';
    const char *p = buf.data();
    Banner b;
    if (parse_banner(path, p, b, err))
        return 1;
    if (!next_int(p, rows) || !next_int(p, cols))
    {
        err = path + ": size line not found";
        return 1;
    }
    return 0;
}

static int read_crs_unguarded(std::string const &path, CsrHost &out, std::string &err,
             long long row_begin, long long row_end)
{
    std::vector<char> buf;
    if (slurp(path, buf, err))
        return 1;
    const char *p = buf.data();
    Banner b;
    if (parse_banner(path, p, b, err))
        return 1;
    if (!b.coordinate)
    {
        err = path + ": a sparse operator needs the coordinate format";
        return 1;
    }
    long long m = 0, n = 0, nz = 0;
    if (!next_int(p, m) || !next_int(p, n) || !next_int(p, nz) || m < 0 || n < 0 || nz < 0)
    {
        err = path + ": bad size line";
        return 1;
    }
    if (m > 2147483647LL || n > 2147483647LL)
    {
        err = path + ": dimensions exceed the int32 index range";
        return 1;
    }
    if (row_end < 0)
        row_end = m;
    if (row_begin < 0 || row_begin > row_end || row_end > m)
    {
        err = path + ": row range outside the matrix";
        return 1;
    }
    const long long local = row_end - row_begin;

    // an entry takes at least four bytes ("1 1\n"): a size line that promises more
    // than the file can hold is corrupt, and must not drive an allocation
    if (nz > (long long)buf.size())
    {
        err = path + ": the size line promises more entries than the file holds";
        return 1;
    }
    std::vector<Triple> entries;
    entries.reserve((size_t)(nz * (b.symmetry ? 2 : 1)));
    for (long long e = 0; e < nz; ++e)
    {
        long long i = 0, j = 0;
        double v = 1.0;
        if (!next_int(p, i) || !next_int(p, j) || (!b.pattern && !next_double(p, v)))
        {
            std::ostringstream os;
            os << path << ": entry " << e + 1 << " of " << nz << " is malformed or missing";
            err = os.str();
            return 1;
        }
        if (i < 1 || i > m || j < 1 || j > n)
        {
            std::ostringstream os;
            os << path << ": entry " << e + 1 << " (" << i << ", " << j << ") is outside the "
               << m << " x " << n << " matrix";
            err = os.str();
            return 1;
        }
        --i;
        --j;
        if (i >= row_begin && i < row_end)
            entries.push_back(Triple{(int)(i - row_begin), (int)j, v});
        if (b.symmetry && i != j && j >= row_begin && j < row_end)
        {
            if (j >= m || i >= n)
            {
                err = path + ": symmetric storage on a non-square matrix";
                return 1;
            }
            entries.push_back(Triple{(int)(j - row_begin), (int)i, b.symmetry == 2 ? -v : v});
        }
    }
    if (entries.size() >= (size_t)2147483647)
    {
        err = path + ": more than 2^31 - 1 entries in one row block";
        return 1;
    }

    // stable sort: entries given twice are added in file order
    std::stable_sort(entries.begin(), entries.end(), [](Triple const &a, Triple const &c) {
        return a.row != c.row ? a.row < c.row : a.col < c.col;
    });

    out.rows = (int)local;
    out.global_rows = (int)m;
    out.cols = (int)n;
    out.row_begin = (int)row_begin;
    out.rowptr.assign((size_t)local + 1, 0);
    out.colidx.clear();
    out.vals.clear();
    out.colidx.reserve(entries.size());
    out.vals.reserve(entries.size());
    for (size_t e = 0; e < entries.size(); ++e)
    {
        if (e > 0 && entries[e].row == entries[e - 1].row && entries[e].col == entries[e - 1].col)
        {
            out.vals.back() += entries[e].val;
            continue;
        }
        out.colidx.push_back(entries[e].col);
        out.vals.push_back(entries[e].val);
        out.rowptr[(size_t)entries[e].row + 1]++;
    }
    for (long long r = 0; r < local; ++r)
        out.rowptr[(size_t)r + 1] += out.rowptr[(size_t)r];
    return 0;
}

static int read_dense_unguarded(std::string const &path, int &rows, int &cols, std::vector<double> &data,
               std::string &err)
{
    std::vector<char> buf;
    if (slurp(path, buf, err))
        return 1;
    const char *p = buf.data();
    Banner b;
    if (parse_banner(path, p, b, err))
        return 1;
    long long m = 0, n = 0;
    if (!next_int(p, m) || !next_int(p, n) || m < 0 || n < 0 || m > 2147483647LL || n > 2147483647LL)
    {
        err = path + ": bad size line";
        return 1;
    }
    // array format: every value takes at least two bytes; a coordinate file is
    // accepted for convenience only, for blocks that fit a sane dense size
    if ((!b.coordinate && (m > 0 && n > (long long)buf.size() / m)) ||
        (b.coordinate && m > 0 && n > (1LL << 28) / m))
    {
        err = path + ": the size line promises more values than the file holds";
        return 1;
    }
    // symmetric / skew-symmetric storage mirrors (i, j) into (j, i): only a square
    // block has room for that (read_crs rejects the same case)
    if (b.symmetry && m != n)
    {
        err = path + ": symmetric storage on a non-square matrix";
        return 1;
    }
    rows = (int)m;
    cols = (int)n;
    data.assign((size_t)m * (size_t)n, 0.0);
    if (b.coordinate)
    {
        long long nz = 0;
        if (!next_int(p, nz))
        {
            err = path + ": bad size line";
            return 1;
        }
        for (long long e = 0; e < nz; ++e)
        {
            long long i = 0, j = 0;
            double v = 1.0;
            if (!next_int(p, i) || !next_int(p, j) || (!b.pattern && !next_double(p, v)) ||
                i < 1 || i > m || j < 1 || j > n)
            {
                err = path + ": malformed entry";
                return 1;
            }
            data[(size_t)(i - 1) + (size_t)(j - 1) * (size_t)m] += v;
            if (b.symmetry && i != j)
                data[(size_t)(j - 1) + (size_t)(i - 1) * (size_t)m] += b.symmetry == 2 ? -v : v;
        }
        return 0;
    }
    if (b.symmetry == 0)
    {
        for (size_t e = 0; e < data.size(); ++e)
            if (!next_double(p, data[e]))
            {
                std::ostringstream os;
                os << path << ": " << data.size() << " values expected, " << e << " found";
                err = os.str();
                return 1;
            }
        return 0;
    }
    // packed lower triangle, column by column (skew-symmetric: strictly lower)
    for (long long j = 0; j < n; ++j)
        for (long long i = j + (b.symmetry == 2 ? 1 : 0); i < m; ++i)
        {
            double v = 0.0;
            if (!next_double(p, v))
            {
                err = path + ": too few values";
                return 1;
            }
            data[(size_t)i + (size_t)j * (size_t)m] = v;
            if (i != j)
                data[(size_t)j + (size_t)i * (size_t)m] = b.symmetry == 2 ? -v : v;
        }
    return 0;
}

// No exception leaves the readers (the reference's error style is an int code,
// src/LyapunovMacros.hpp:7-26): an allocation a corrupt file still manages to
// provoke becomes an error message.
template<class F>
static int guarded(std::string const &path, std::string &err, F body)
{
    try
    {
        return body();
    }
    catch (std::exception const &e)
    {
        err = path + ": " + e.what();
        return 1;
    }
}

int read_size(std::string const &path, long long &rows, long long &cols, std::string &err)
{
    return guarded(path, err, [&] { return read_size_unguarded(path, rows, cols, err); });
}

int read_crs(std::string const &path, CsrHost &out, std::string &err, long long row_begin, long long row_end)
{
    return guarded(path, err, [&] { return read_crs_unguarded(path, out, err, row_begin, row_end); });
}

int read_dense(std::string const &path, int &rows, int &cols, std::vector<double> &data, std::string &err)
{
    return guarded(path, err, [&] { return read_dense_unguarded(path, rows, cols, data, err); });
}

int write_dense(std::string const &path, int rows, int cols, const double *data, long long ld,
                std::string const &comment, std::string &err)
{
    FILE *f = std::fopen(path.c_str(), "wb");
    if (!f)
    {
        err = "cannot create " + path + ": " + std::strerror(errno);
        return 1;
    }
    std::vector<char> iobuf(1 << 20);
    std::setvbuf(f, iobuf.data(), _IOFBF, iobuf.size());
    std::fprintf(f, "%%%%MatrixMarket matrix array real general\n");
    if (!comment.empty())
        std::fprintf(f, "%% %s\n", comment.c_str());
    std::fprintf(f, "%d %d\n", rows, cols);
    for (int j = 0; j < cols; ++j)
        for (int i = 0; i < rows; ++i)
            std::fprintf(f, "%.17g\n", data[(size_t)i + (size_t)j * (size_t)ld]);
    int bad = std::ferror(f);
    if (std::fclose(f) || bad)
    {
        err = "write error on " + path;
        return 1;
    }
    return 0;
}

int write_crs(std::string const &path, CsrHost const &A, std::string &err)
{
    FILE *f = std::fopen(path.c_str(), "wb");
    if (!f)
    {
        err = "cannot create " + path + ": " + std::strerror(errno);
        return 1;
    }
    std::vector<char> iobuf(1 << 20);
    std::setvbuf(f, iobuf.data(), _IOFBF, iobuf.size());
    std::fprintf(f, "%%%%MatrixMarket matrix coordinate real general\n");
    std::fprintf(f, "%d %d %zu\n", A.global_rows, A.cols, A.vals.size());
    for (int r = 0; r < A.rows; ++r)
        for (int e = A.rowptr[r]; e < A.rowptr[r + 1]; ++e)
            std::fprintf(f, "%d %d %.17g\n", r + A.row_begin + 1, A.colidx[e] + 1, A.vals[e]);
    int bad = std::ferror(f);
    if (std::fclose(f) || bad)
    {
        err = "write error on " + path;
        return 1;
    }
    return 0;
}

bool is_diagonal(CsrHost const &A)
{
    for (int r = 0; r < A.rows; ++r)
        for (int e = A.rowptr[r]; e < A.rowptr[r + 1]; ++e)
            if (A.colidx[e] != r + A.row_begin && A.vals[e] != 0.0)
                return false;
    return true;
}

bool is_identity(CsrHost const &A)
{
    if (A.global_rows != A.cols || !is_diagonal(A))
        return false;
    for (int r = 0; r < A.rows; ++r)
    {
        double d = 0.0;
        for (int e = A.rowptr[r]; e < A.rowptr[r + 1]; ++e)
            if (A.colidx[e] == r + A.row_begin)
                d += A.vals[e];
        if (d != 1.0)
            return false;
    }
    return true;
}

long long zero_diagonal_rows(CsrHost const &A)
{
    long long count = 0;
    for (int r = 0; r < A.rows; ++r)
    {
        double d = 0.0;
        for (int e = A.rowptr[r]; e < A.rowptr[r + 1]; ++e)
            if (A.colidx[e] == r + A.row_begin)
                d += A.vals[e];
        if (d == 0.0)
            ++count;
    }
    return count;
}

}
}
