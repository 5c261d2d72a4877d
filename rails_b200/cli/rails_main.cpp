// rails_cuda_main -- the command-line driver of the B200 backend
// (SURVEY.md §8(f) N1 + N2).
//
// Same job as the reference's only executable (/root/reference/src/main.cpp):
// read A.mtx, B.mtx and M.mtx from the working directory, read the solver
// parameters from the file named on the command line, solve
//     A X M' + M X A' + B B' = 0,   X = V T V',
// write V.mtx and T.mtx, print the dominant eigenvalues of X next to their
// share of trace(X), and dump the timing profile.  Differences, all forced by
// what is in scope:
//   * the reference always solves on the Schur complement of the rows where M is
//     zero (src/main.cpp:75-98).  Here that reduction (CudaSchurOperator, one
//     GPU) is used when M has zero diagonal rows and A itself otherwise; V.mtx
//     then holds the PROLONGATED basis (all m rows), so that X = V T V' is the
//     solution of the original equation and the eigenvalues printed are those
//     the reference gets from SchurOperator::Apply with SetSolution;
//   * a non-identity (reduced) M selects the generalised branch ("Use mass
//     matrix", matlab/RAILSsolver.m:368-392), an identity or absent M the
//     standard one; "Projection method" > 1 uses S^-1 x = (A^-1 [0; x])(idx2)
//     as opts.Ainv (matlab/RAILSschur.m:60-66);
//   * the eigenpairs of X come from the k x k matrix T and one V*C product
//     (SolutionTools.hpp) instead of Anasazi;
//   * one process drives one GPU.  Started N times with RANK / WORLD_SIZE /
//     LOCAL_RANK in the environment (what torchrun --no-python or mpirun
//     wrappers export) the ranks partition the rows like Epetra_Map and meet
//     through a rendezvous file for the NCCL id.
//
// Usage: rails_cuda_main [params.xml|params.ini] [--dir DIR] [--A f] [--B f]
//        [--M f] [--V f] [--T f] [--only-eigenvalues] [--write-eigs]
//        [--b-operator] [--device N] [--seed S]
#include "CudaFusedLanczos.hpp"
#include "CudaMultiVectorWrapper.hpp"
#include "CudaOperatorWrapper.hpp"
#include "CudaSchurOperator.hpp"
#include "HostDenseMatrix.hpp"
#include "MatrixMarket.hpp"
#include "ParameterFile.hpp"
#include "RailsSolver.hpp"
#include "SolutionTools.hpp"

#include <chrono>
#include <ctime>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iomanip>
#include <iostream>
#include <string>
#include <thread>
#include <sys/stat.h>
#include <unistd.h>

using namespace RAILS;

extern "C" void rails_host_limit_blas_threads(void);

namespace
{

typedef b200::Solver<CudaOperatorWrapper, CudaMultiVectorWrapper, HostDenseMatrix> CudaSolver;

struct Options
{
    std::string params;
    std::string dir = ".";
    std::string A = "A.mtx", B = "B.mtx", M = "M.mtx", V = "V.mtx", T = "T.mtx";
    bool only_eigenvalues = false;
    bool write_eigs = false;
    bool b_operator = false;
    int device = -1;
    unsigned seed = 1;
};

int env_int(const char *name, int def)
{
    const char *e = std::getenv(name);
    return e && *e ? std::atoi(e) : def;
}

std::string join(std::string const &dir, std::string const &file)
{
    if (!file.empty() && file[0] == '/')
        return file;
    return dir + "/" + file;
}

bool exists(std::string const &path)
{
    return access(path.c_str(), R_OK) == 0;
}

int usage(const char *argv0)
{
    std::cerr << "usage: " << argv0 << " [params.xml|params.ini] [--dir DIR] [--A f] [--B f] [--M f]\n"
              << "       [--V f] [--T f] [--only-eigenvalues] [--write-eigs] [--b-operator]\n"
              << "       [--device N] [--seed S]" << std::endl;
    return 64;
}

// NCCL id through a file: rank 0 writes it under a temporary name and renames
// it into place, the others poll for it
int exchange_id(std::string const &path, int rank, char id[RAILS_NCCL_ID_BYTES])
{
    if (rank == 0)
    {
        if (rails_cuda_comm_unique_id(id))
            return 1;
        std::string tmp = path + ".tmp";
        FILE *f = std::fopen(tmp.c_str(), "wb");
        if (!f)
            return 1;
        std::fwrite(id, 1, RAILS_NCCL_ID_BYTES, f);
        std::fclose(f);
        return std::rename(tmp.c_str(), path.c_str());
    }
    // a file left behind by a run that died before rank 0 could remove it is older
    // than this launch: ignore anything written more than two minutes before now
    const time_t oldest = std::time(nullptr) - 120;
    for (int tries = 0; tries < 6000; ++tries)
    {
        struct stat st;
        FILE *f = stat(path.c_str(), &st) == 0 && st.st_mtime >= oldest ? std::fopen(path.c_str(), "rb") : nullptr;
        if (f)
        {
            size_t got = std::fread(id, 1, RAILS_NCCL_ID_BYTES, f);
            std::fclose(f);
            if (got == RAILS_NCCL_ID_BYTES)
                return 0;
        }
        std::this_thread::sleep_for(std::chrono::milliseconds(10));
    }
    return 1;
}

// cross-rank OR of a local flag (one tiny reduction over the distributed map)
bool any_rank(bool local, int global_rows)
{
    if (rails_cuda_comm_size() == 1)
        return local;
    CudaMultiVectorWrapper x(global_rows, 1);
    x = local ? 1.0 : 0.0;
    return x.norm() > 0.0;
}

// the reduction doubles as a barrier: no rank leaves before all have arrived
void barrier(int global_rows)
{
    any_rank(false, global_rows);
}

// V (distributed) -> one array-format file written by rank 0
int write_distributed(std::string const &path, CudaMultiVectorWrapper const &V, int rank, int nranks,
                      std::string const &comment, std::string &err)
{
    const int m = V.M(), k = V.N();
    const long long local = V.local_rows();
    std::vector<double> mine((size_t)std::max<long long>(local, 1) * (size_t)std::max(k, 1));
    if (k > 0 && V.to_host(mine.data(), std::max<long long>(local, 1)))
    {
        err = "download of " + path + " failed";
        return 1;
    }
    if (nranks == 1)
        return io::write_dense(path, m, k, mine.data(), std::max<long long>(local, 1), comment, err);

    std::string part = path + ".part" + std::to_string(rank);
    FILE *f = std::fopen(part.c_str(), "wb");
    if (!f)
    {
        err = "cannot create " + part;
        return 1;
    }
    std::fwrite(mine.data(), sizeof(double), (size_t)local * (size_t)k, f);
    std::fclose(f);
    barrier(m);
    int rc = 0;
    if (rank == 0)
    {
        std::vector<double> all((size_t)m * (size_t)std::max(k, 1));
        for (int r = 0; r < nranks && !rc; ++r)
        {
            long long b = 0, e = 0;
            rails_partition_rows(m, nranks, r, &b, &e);
            std::string p = path + ".part" + std::to_string(r);
            FILE *g = std::fopen(p.c_str(), "rb");
            if (!g)
            {
                err = "cannot open " + p;
                rc = 1;
                break;
            }
            std::vector<double> blk((size_t)(e - b) * (size_t)std::max(k, 1));
            size_t want = (size_t)(e - b) * (size_t)k;
            if (std::fread(blk.data(), sizeof(double), want, g) != want)
            {
                err = "short read on " + p;
                rc = 1;
            }
            std::fclose(g);
            std::remove(p.c_str());
            for (int j = 0; j < k && !rc; ++j)
                std::memcpy(&all[(size_t)b + (size_t)j * (size_t)m], &blk[(size_t)j * (size_t)(e - b)],
                            sizeof(double) * (size_t)(e - b));
        }
        if (!rc)
            rc = io::write_dense(path, m, k, all.data(), m, comment, err);
    }
    barrier(m);
    return rc;
}

}

int main(int argc, char *argv[])
{
    Options opt;
    for (int i = 1; i < argc; ++i)
    {
        std::string a = argv[i];
        auto value = [&](std::string &dst) {
            if (i + 1 >= argc)
                return false;
            dst = argv[++i];
            return true;
        };
        std::string tmp;
        if (a == "--dir") { if (!value(opt.dir)) return usage(argv[0]); }
        else if (a == "--A") { if (!value(opt.A)) return usage(argv[0]); }
        else if (a == "--B") { if (!value(opt.B)) return usage(argv[0]); }
        else if (a == "--M") { if (!value(opt.M)) return usage(argv[0]); }
        else if (a == "--V") { if (!value(opt.V)) return usage(argv[0]); }
        else if (a == "--T") { if (!value(opt.T)) return usage(argv[0]); }
        else if (a == "--device") { if (!value(tmp)) return usage(argv[0]); opt.device = std::atoi(tmp.c_str()); }
        else if (a == "--seed") { if (!value(tmp)) return usage(argv[0]); opt.seed = (unsigned)std::strtoul(tmp.c_str(), nullptr, 10); }
        else if (a == "--only-eigenvalues") opt.only_eigenvalues = true;
        else if (a == "--write-eigs") opt.write_eigs = true;
        else if (a == "--b-operator") opt.b_operator = true;
        else if (a == "-h" || a == "--help") { usage(argv[0]); return 0; }
        else if (!a.empty() && a[0] == '-') return usage(argv[0]);
        else opt.params = a;
    }

    const int rank = env_int("RANK", 0), nranks = std::max(env_int("WORLD_SIZE", 1), 1);
    const int device = opt.device >= 0 ? opt.device : env_int("LOCAL_RANK", 0);
    const bool root = rank == 0;
    std::string err;

    ParameterTree params("RAILS");
    if (!opt.params.empty() && ParameterTree::from_file(opt.params, params, err))
    {
        std::cerr << err << std::endl;
        return 2;
    }

    if (rails_cuda_init(device))
    {
        std::cerr << "rails_cuda_init(" << device << ") failed: " << rails_cuda_last_error() << std::endl;
        return 3;
    }
    rails_host_limit_blas_threads();
    if (nranks > 1)
    {
        const char *rv = std::getenv("RAILS_RENDEZVOUS");
        std::string file = rv && *rv ? std::string(rv)
                                     : join(opt.dir, ".rails_nccl_id." + std::to_string(env_int("MASTER_PORT", 0)));
        char id[RAILS_NCCL_ID_BYTES];
        if (exchange_id(file, rank, id) || rails_cuda_comm_init(rank, nranks, id))
        {
            std::cerr << "rank " << rank << ": communicator setup failed: " << rails_cuda_last_error()
                      << std::endl;
            if (rank == 0)
                std::remove(file.c_str());
            return 3;
        }
        // the communicator exists on every rank, so every rank has read the id
        if (rank == 0)
            std::remove(file.c_str());
    }

    std::map<std::string, double> profile;
    auto tick = [] { return std::chrono::steady_clock::now(); };
    auto since = [](std::chrono::steady_clock::time_point t0) {
        return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    };

    if (root)
        std::cout << "Loading matrices" << std::endl;
    auto t_load = tick();

    long long m = 0, ncols = 0;
    if (io::read_size(join(opt.dir, opt.A), m, ncols, err))
    {
        std::cerr << err << std::endl;
        return 2;
    }
    if (m != ncols)
    {
        std::cerr << opt.A << ": A must be square, it is " << m << " x " << ncols << std::endl;
        return 2;
    }
    long long r0 = 0, r1 = m;
    rails_partition_rows(m, nranks, rank, &r0, &r1);

    io::CsrHost A, B, M;
    if (io::read_crs(join(opt.dir, opt.A), A, err, r0, r1) ||
        io::read_crs(join(opt.dir, opt.B), B, err, r0, r1))
    {
        std::cerr << err << std::endl;
        return 2;
    }
    if (B.global_rows != m)
    {
        std::cerr << opt.B << ": B has " << B.global_rows << " rows, A has " << m << std::endl;
        return 2;
    }
    bool have_M = exists(join(opt.dir, opt.M));
    if (have_M)
    {
        if (io::read_crs(join(opt.dir, opt.M), M, err, r0, r1))
        {
            std::cerr << err << std::endl;
            return 2;
        }
        if (M.global_rows != m || M.cols != m)
        {
            std::cerr << opt.M << ": M must be " << m << " x " << m << std::endl;
            return 2;
        }
    }
    profile["Load matrices"] = since(t_load);

    ParameterTree &solver_params = params.sublist("Lyapunov Solver");
    bool mass = false;
    bool singular = false;
    if (have_M)
    {
        singular = any_rank(io::zero_diagonal_rows(M) > 0, (int)m);
        if (singular && (nranks > 1 || opt.b_operator))
        {
            if (root)
                std::cerr << "M has rows with a zero diagonal: the Schur complement reduction runs on one\n"
                             "GPU with a dense B (no --b-operator)." << std::endl;
            return 4;
        }
        mass = !singular && any_rank(!io::is_identity(M), (int)m);
    }
    CudaSchurOperator schur;
    if (singular)
    {
        if (root)
            std::cout << "Computing Schur complement" << std::endl;
        auto t_schur = tick();
        std::vector<double> mdiag((size_t)m, 0.0);
        for (int r = 0; r < M.rows; ++r)
            for (int e = M.rowptr[r]; e < M.rowptr[r + 1]; ++e)
                if (M.colidx[e] == r)
                    mdiag[r] += M.vals[e];
        schur = CudaSchurOperator((int)m, A.rowptr.data(), A.colidx.data(), A.vals.data(), mdiag.data(),
                                  solver_params.get("Schur zero tolerance", 1e-12),
                                  solver_params.get("Schur scale", 1.0));
        int src = schur.Compute();
        if (src)
        {
            std::cerr << "Schur complement set-up failed (" << src << ")" << std::endl;
            return 4;
        }
        mass = !schur.mass_is_identity();
        profile["Compute Schur complement"] = since(t_schur);
        std::cout << "Singular part: " << schur.n1() << " rows, Schur complement: " << schur.n2()
                  << " rows, nnz(L+U) of A11: " << schur.factor_nnz() << ", levels "
                  << schur.factor_levels(0) << "/" << schur.factor_levels(1) << ", "
                  << schur.launches_per_solve() << " launches per solve" << std::endl;
    }
    if (mass && !solver_params.isParameter("Use mass matrix"))
        solver_params.set("Use mass matrix", true);
    mass = solver_params.get("Use mass matrix", false);
    if (singular && !mass)
        solver_params.set("Use mass matrix", false);
    if (mass && !have_M)
    {
        std::cerr << "\"Use mass matrix\" is set but " << opt.M << " does not exist" << std::endl;
        return 2;
    }

    if (root)
        std::cout << "Creating solver" << std::endl;
    auto t_setup = tick();
    CudaOperatorWrapper Aop;
    if (singular)
        Aop = schur.schur_operator();
    else
        Aop = CudaOperatorWrapper(A.rows, (int)m, (int)m, (int)r0, A.rowptr.data(), A.colidx.data(),
                                  A.vals.data());
    if (!Aop.valid())
        return 3;
    CudaOperatorWrapper Mop = Aop;
    if (singular)
        Mop = schur.mass_operator();
    else if (have_M)
    {
        Mop = CudaOperatorWrapper(M.rows, (int)m, (int)m, (int)r0, M.rowptr.data(), M.colidx.data(),
                                  M.vals.data());
        if (!Mop.device_handle())
            return 3;
    }

    // B: dense multivector by default (the fused Lanczos chain wants one), or
    // the sparse operator itself like the reference's main (src/main.cpp:96-98)
    const int n_b = B.cols;
    CudaMultiVectorWrapper Bmv;
    CudaOperatorWrapper Bop;
    if (opt.b_operator)
    {
        Bop = CudaOperatorWrapper(B.rows, (int)m, n_b, (int)r0, B.rowptr.data(), B.colidx.data(),
                                  B.vals.data());
        if (!Bop.device_handle())
            return 3;
    }
    else
    {
        const long long local = r1 - r0;
        std::vector<double> dense((size_t)std::max<long long>(local, 1) * (size_t)std::max(n_b, 1), 0.0);
        for (int r = 0; r < B.rows; ++r)
            for (int e = B.rowptr[r]; e < B.rowptr[r + 1]; ++e)
                dense[(size_t)r + (size_t)B.colidx[e] * (size_t)local] = B.vals[e];
        Bmv = CudaMultiVectorWrapper((int)m, n_b);
        if (rails_mv_upload(Bmv.device_handle(), 0, n_b, dense.data(), std::max<long long>(local, 1)))
        {
            std::cerr << "upload of B failed: " << rails_cuda_last_error() << std::endl;
            return 3;
        }
    }
    if (singular)
        Bmv = schur.restrict(Bmv);          // BS (matlab/RAILSschur.m:43-49)
    profile["Create operators"] = since(t_setup);

    CudaMultiVectorWrapper V;
    HostDenseMatrix T;
    int ret = 0;

    const bool warm = solver_params.get("Restart from solution", false);
    if (opt.only_eigenvalues || warm)
    {
        // V.mtx / T.mtx from a previous run (src/main.cpp:155-164)
        int vr = 0, vc = 0;
        std::vector<double> data;
        if (io::read_dense(join(opt.dir, opt.V), vr, vc, data, err))
        {
            std::cerr << err << std::endl;
            return 2;
        }
        if (vr != m)
        {
            std::cerr << opt.V << ": " << vr << " rows, expected " << m << std::endl;
            return 2;
        }
        V = CudaMultiVectorWrapper::from_host((int)m, vc, data.data(), m);
        if (singular && !opt.only_eigenvalues)
            V = schur.select_regular(V);    // the reduced basis sits in the rows idx2
        if (opt.only_eigenvalues)
        {
            int tr = 0, tc = 0;
            std::vector<double> tdata;
            if (io::read_dense(join(opt.dir, opt.T), tr, tc, tdata, err))
            {
                std::cerr << err << std::endl;
                return 2;
            }
            if (tr != vc || tc != vc)
            {
                std::cerr << opt.T << ": " << tr << " x " << tc << ", expected " << vc << " x " << vc
                          << std::endl;
                return 2;
            }
            T = HostDenseMatrix(tr, tc);
            for (int j = 0; j < tc; ++j)
                for (int i = 0; i < tr; ++i)
                    T(i, j) = tdata[(size_t)i + (size_t)j * (size_t)tr];
        }
    }
    else
        V = CudaMultiVectorWrapper(singular ? schur.n2() : (int)m, 1);

    if (!opt.only_eigenvalues)
    {
        std::srand(opt.seed);
        int prc = 0;
        std::map<std::string, double> sections;
        std::vector<b200::IterationRecord> history;
        auto run = [&](CudaSolver &solver) {
            prc = solver.set_parameters(solver_params);
            if (prc)
                return;
            solver.set_verbose(root && solver_params.get("Verbose", true));
            if (singular && solver_params.get("Projection method", 1.0) > 1.0 + 1e-8)
                solver.set_inverse(schur.inverse_operator());
            if (root)
                std::cout << "Performing solve" << std::endl;
            if (root && singular)
                std::cout << "Amount of matrix-vector products before the solve: " << schur.GetMVPs()
                          << std::endl;
            auto t_solve = tick();
            ret = solver.solve(V, T);
            rails_cuda_sync();
            profile["Solver"] = since(t_solve);
            sections = solver.sections();
            history = solver.history();
        };
        if (opt.b_operator)
        {
            CudaSolver solver(Aop, Bop, Mop);
            run(solver);
        }
        else
        {
            CudaSolver solver(Aop, Bmv, Mop);
            run(solver);
        }
        if (prc)
            return 2;
        for (auto const &kv : sections)
            profile["Solver: " + kv.first] = kv.second;
        if (root && !history.empty())
            std::cout << "Iterations: " << history.size() << ", final relative residual estimate: "
                      << history.back().relative << ", space size: " << V.N() << std::endl;
        if (singular)
        {
            std::cout << "Amount of matrix-vector products after the solve: " << schur.GetMVPs() << std::endl;
            auto t_pro = tick();
            V = schur.prolongate(V);        // Vtrans (matlab/RAILSschur.m:68): X = V T V' on all m rows
            rails_cuda_sync();
            profile["Prolongate solution"] = since(t_pro);
        }

        auto t_write = tick();
        if (write_distributed(join(opt.dir, opt.V), V, rank, nranks, "RAILS solution factor V", err))
        {
            std::cerr << err << std::endl;
            return 5;
        }
        if (root)
        {
            const int k = T.M();
            std::vector<double> t((size_t)std::max(k, 1) * (size_t)std::max(k, 1));
            for (int j = 0; j < k; ++j)
                for (int i = 0; i < k; ++i)
                    t[(size_t)i + (size_t)j * (size_t)k] = T(i, j);
            if (io::write_dense(join(opt.dir, opt.T), k, k, t.data(), std::max(k, 1),
                                "RAILS solution factor T", err))
            {
                std::cerr << err << std::endl;
                return 5;
            }
        }
        profile["Write solution"] = since(t_write);
    }

    // ---- dominant eigenpairs of X and their share of the trace --------------
    ParameterTree &eig_params = params.sublist("Eigenvalue Solver");
    const int num = eig_params.get("Number of Eigenvalues", 10);
    const double eig_tol = eig_params.get("Convergence Tolerance", 1e-6);

    auto t_eigs = tick();
    CudaMultiVectorWrapper eigenvectors;
    HostDenseMatrix eigenvalues(1, 1);
    eigenvalues.resize(0, 1);
    int erc = b200::solution_eigs(V, T, eigenvectors, eigenvalues, num, eig_tol);
    rails_cuda_sync();
    profile["Compute eigenvalues"] = since(t_eigs);

    auto t_trace = tick();
    const double trace = b200::solution_trace(V, T);
    if (root && !erc)
    {
        std::cout << "Eigenvalues of X and their share of trace(X) = " << std::setprecision(12) << trace
                  << std::endl;
        for (int i = 0; i < eigenvalues.M(); i++)
            std::cout << std::setw(20) << eigenvalues(i) << std::setw(20) << eigenvalues(i) / trace
                      << std::endl;
    }
    profile["Compute trace"] = since(t_trace);

    if (opt.write_eigs && !erc)
    {
        if (write_distributed(join(opt.dir, "eigenvectors.mtx"), eigenvectors, rank, nranks,
                              "dominant eigenvectors of X = V T V'", err))
        {
            std::cerr << err << std::endl;
            return 5;
        }
        if (root)
        {
            const int r = eigenvalues.M();
            std::vector<double> ev((size_t)std::max(r, 1));
            for (int i = 0; i < r; ++i)
                ev[i] = eigenvalues(i);
            if (io::write_dense(join(opt.dir, "eigenvalues.mtx"), r, 1, ev.data(), std::max(r, 1),
                                "dominant eigenvalues of X = V T V'", err))
            {
                std::cerr << err << std::endl;
                return 5;
            }
        }
    }

    // timing profile (RAILS_SAVE_PROFILES(""), src/main.cpp:172-173)
    if (root)
    {
        std::cout << "\nProfile (wall-clock seconds on rank 0, " << nranks << " GPU(s), "
                  << rails_cuda_kernel_launches() << " kernel launches)" << std::endl;
        for (auto const &kv : profile)
            std::cout << "  " << std::left << std::setw(36) << kv.first << std::right << std::setw(14)
                      << std::fixed << std::setprecision(6) << kv.second << std::endl;
    }

    // release device objects before the context goes away
    eigenvectors = CudaMultiVectorWrapper();
    V = CudaMultiVectorWrapper();
    Bmv = CudaMultiVectorWrapper();
    Bop = CudaOperatorWrapper();
    Aop = CudaOperatorWrapper();
    Mop = CudaOperatorWrapper();
    schur = CudaSchurOperator();
    rails_cuda_finalize();

    // the solver's own code: 0 converged, -1 stopped unconverged, 1 fell through
    return ret == 0 ? (erc ? 6 : 0) : (ret < 0 ? 10 : 10 + ret);
}
