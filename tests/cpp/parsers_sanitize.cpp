// Driver of tests/test_host_sanitizers.py: the MatrixMarket and parameter-file
// readers (rails_b200/host/MatrixMarket.cpp, ParameterFile.cpp) fed with valid,
// truncated and corrupted files under AddressSanitizer + UBSan.  The readers must
// return an error code or a consistent result -- never read or write out of bounds.
#include "MatrixMarket.hpp"
#include "ParameterFile.hpp"

#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <string>
#include <vector>

using namespace RAILS;

static void write_file(std::string const &path, std::string const &text)
{
    std::ofstream f(path.c_str(), std::ios::binary);
    f << text;
}

static int check_csr(io::CsrHost const &A)
{
    if ((int)A.rowptr.size() != A.rows + 1) return 1;
    if (A.rowptr[0] != 0 || A.rowptr[A.rows] != (int)A.colidx.size() || A.colidx.size() != A.vals.size()) return 1;
    for (int i = 0; i < A.rows; ++i)
        if (A.rowptr[i + 1] < A.rowptr[i]) return 1;
    for (size_t p = 0; p < A.colidx.size(); ++p)
        if (A.colidx[p] < 0 || A.colidx[p] >= A.cols) return 1;
    return 0;
}

int main(int argc, char **argv)
{
    const std::string dir = argc > 1 ? argv[1] : ".";
    const std::string good_mm =
        "%%MatrixMarket matrix coordinate real general\n% comment\n4 4 6\n1 1 2.5\n2 2 -1\n3 1 4\n4 4 1e-3\n2 3 7\n2 3 1\n";
    const std::string sym_mm = "%%MatrixMarket matrix coordinate real symmetric\n3 3 3\n1 1 1\n2 1 5\n3 3 2\n";
    const std::string arr_mm = "%%MatrixMarket matrix array real general\n3 2\n1\n2\n3\n4\n5\n6\n";
    const std::string good_xml =
        "<ParameterList name=\"RAILS\">\n <ParameterList name=\"Lyapunov Solver\">\n"
        "  <Parameter name=\"Tolerance\" type=\"double\" value=\"1e-6\"/>\n"
        "  <Parameter name=\"Expand size\" type=\"int\" value=\"2\"/>\n </ParameterList>\n</ParameterList>\n";
    const std::string good_ini = "[Lyapunov Solver]\nTolerance = 1e-5\nRestart from solution = true\n; note\n[Eigenvalue Solver]\nNumber of Eigenvalues = 4\n";

    int bad = 0, parsed = 0, rejected = 0;
    std::srand(7);
    std::vector<std::string> seeds = {good_mm, sym_mm, arr_mm, good_xml, good_ini,
        "%%MatrixMarket matrix coordinate real general\n2 2 1\n3 1 1.0\n",          // row out of range
        "%%MatrixMarket matrix coordinate real general\n2 2 5\n1 1 1.0\n",          // fewer entries than promised
        "%%MatrixMarket matrix coordinate pattern general\n2 2 2\n1 2\n2 1\n",
        "%%MatrixMarket matrix coordinate real general\n-1 4 2\n",
        // symmetric storage on a non-square size line: the mirror write would leave the buffer
        "%%MatrixMarket matrix array real symmetric\n4 2\n1\n2\n3\n4\n5\n6\n7\n",
        "%%MatrixMarket matrix coordinate real symmetric\n4 2 2\n4 1 1.0\n3 2 2.0\n",
        "%%MatrixMarket matrix array real skew-symmetric\n5 3\n1\n2\n3\n4\n5\n6\n7\n",
        "%%MatrixMarket matrix coordinate real general\n2000000000 2000000000 1\n1 1 1\n",
        "", "\n\n\n", "<ParameterList", "[unterminated\nkey", "<Parameter name=\"x\" type=\"int\" value=\"abc\"/>"};
    // mutations of the valid files: truncations, byte flips, duplicated chunks
    for (int s = 0; s < 5; ++s)
        for (int m = 0; m < 120; ++m)
        {
            std::string t = seeds[s];
            const int kind = std::rand() % 3;
            if (kind == 0) t = t.substr(0, std::rand() % (t.size() + 1));
            else if (kind == 1) for (int k = 0; k < 3; ++k) t[std::rand() % t.size()] = (char)(std::rand() % 256);
            else { size_t a = std::rand() % t.size(); t.insert(a, t.substr(a, std::rand() % 40)); }
            seeds.push_back(t);
        }
    for (size_t i = 0; i < seeds.size(); ++i)
    {
        const std::string path = dir + "/case.txt";
        write_file(path, seeds[i]);
        std::string err;
        io::CsrHost A;
        if (io::read_crs(path, A, err) == 0) { ++parsed; bad |= check_csr(A); } else ++rejected;
        io::CsrHost Ablock;
        if (io::read_crs(path, Ablock, err, 1, 3) == 0) bad |= check_csr(Ablock);
        long long r = 0, c = 0;
        io::read_size(path, r, c, err);
        int rows = 0, cols = 0;
        std::vector<double> data;
        if (io::read_dense(path, rows, cols, data, err) == 0)
        {
            ++parsed;
            if (data.size() != (size_t)rows * (size_t)cols) bad = 1;
        }
        ParameterTree tree;
        if (ParameterTree::from_file(path, tree, err) == 0)
        {
            ++parsed;
            (void)tree.sublist("Lyapunov Solver").get("Tolerance", 1e-3);
            (void)tree.sublist("Lyapunov Solver").get("Expand size", 3);
            (void)tree.sublist("Eigenvalue Solver").get("Number of Eigenvalues", 10);
        }
    }
    // the valid files parse to what they say
    {
        std::string err;
        write_file(dir + "/good.mtx", good_mm);
        io::CsrHost A;
        if (io::read_crs(dir + "/good.mtx", A, err) || A.rows != 4 || A.cols != 4 || A.vals.size() != 5) bad = 1;   // the duplicate (2,3) is summed
        write_file(dir + "/good.xml", good_xml);
        ParameterTree t;
        if (ParameterTree::from_file(dir + "/good.xml", t, err) || t.sublist("Lyapunov Solver").get("Expand size", 3) != 2) bad = 1;
        write_file(dir + "/good.ini", good_ini);
        ParameterTree u;
        if (ParameterTree::from_file(dir + "/good.ini", u, err) || !u.sublist("Lyapunov Solver").get("Restart from solution", false)) bad = 1;
    }
    std::printf("%zu files, %d parsed, %d rejected\n", seeds.size(), parsed, rejected);
    std::printf(bad ? "FAILED\n" : "OK\n");
    return bad;
}
