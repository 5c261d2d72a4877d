// The reference's algorithm integration tests (/root/reference/test/LyapunovSolver_test.cpp:118-352:
// dense 20 x 20 and random tridiagonal problems, restart by size, by iterations,
// minimised solution space, warm start) on the CUDA wrappers.
//
// Those test bodies are hard-wired to ONE class that is Operator, MultiVector and
// DenseMatrix at once (`Solver<StlWrapper, StlWrapper, StlWrapper>`, residual formed as
// `A * X * T * X.transpose() + ...` on that class), so they cannot be re-typed through a
// Types<> list like the three Generic* suites.  They are restated here case by case --
// same problem construction through the reference's own StlWrapper (same random
// stream), same parameter lists, same assertions -- and run for BOTH solver loops over
// the CUDA wrappers:
//     RAILS::Solver<...>        the UNMODIFIED reference loop (src/LyapunovSolver.hpp)
//     RAILS::b200::Solver<...>  the product loop (rails_b200/host/RailsSolver.hpp)
// The residual A X T X' + X T X' A' + B B' is formed on the host with StlWrapper
// arithmetic from the downloaded factors and every entry is held to the reference's
// 1e-3 (test/TestHelpers.hpp:4-11).
#include <algorithm>
#include <iostream>
#include <map>
#include <string>
#include <utility>
#include <vector>

#include "CudaMultiVectorWrapper.hpp"
#include "CudaOperatorWrapper.hpp"
#include "HostDenseMatrix.hpp"
#include "RailsSolver.hpp"
#include "CudaFusedLanczos.hpp"

#include "gtest/gtest.h"
#include "test/TestHelpers.hpp"

#include "src/LyapunovSolver.hpp"
#include "src/StlWrapper.hpp"

using namespace RAILS;

namespace
{

// the duck-typed parameter list of test/LyapunovSolver_test.cpp:160-179
class ParameterList
{
    std::map<std::string, double> params_;

public:
    template<typename T>
    T get(std::string const &name, T def)
    {
        auto it = params_.find(name);
        if (it == params_.end())
            return def;
        return it->second;
    }
    template<typename T>
    void set(std::string const &name, T val) { params_[name] = val; }
};

// test/LyapunovSolver_test.cpp:181-200
void get_tridiagonal_problem(int n, StlWrapper &A, StlWrapper &B)
{
    A = StlWrapper(n, n);
    A.random();
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j)
            if (std::abs(i - j) > 1)
                A(i, j) = 0.0;
            else if (i == j)
                A(i, j) *= 3.0;
    B = StlWrapper(n, 1);
    B.random();
    StlWrapper C = B.copy();
    C(n - 1, 0) = 0.0;
    B -= C;
}

CudaOperatorWrapper to_device_operator(StlWrapper const &A)
{
    const int m = A.M(), n = A.N();
    std::vector<int> rowptr(m + 1, 0), colidx;
    std::vector<double> vals;
    for (int i = 0; i < m; ++i)
    {
        for (int j = 0; j < n; ++j)
            if (A(i, j) != 0.0)
            {
                colidx.push_back(j);
                vals.push_back(A(i, j));
            }
        rowptr[i + 1] = (int)colidx.size();
    }
    colidx.push_back(0);
    vals.push_back(0.0);
    return CudaOperatorWrapper(m, n, rowptr.data(), colidx.data(), vals.data());
}

CudaMultiVectorWrapper to_device_multivector(StlWrapper const &B)
{
    std::vector<double> host((size_t)B.M() * B.N());
    for (int j = 0; j < B.N(); ++j)
        for (int i = 0; i < B.M(); ++i)
            host[i + (size_t)j * B.M()] = B(i, j);
    return CudaMultiVectorWrapper::from_host(B.M(), B.N(), host.data(), B.M());
}

StlWrapper to_host(CudaMultiVectorWrapper const &X)
{
    StlWrapper out(X.M(), std::max(X.N(), 1));
    out.resize(X.N());
    std::vector<double> host((size_t)X.M() * std::max(X.N(), 1));
    X.to_host(host.data(), X.M());
    for (int j = 0; j < X.N(); ++j)
        for (int i = 0; i < X.M(); ++i)
            out(i, j) = host[i + (size_t)j * X.M()];
    return out;
}

StlWrapper to_host(HostDenseMatrix const &T)
{
    StlWrapper out(T.M(), T.N());
    for (int j = 0; j < T.N(); ++j)
        for (int i = 0; i < T.M(); ++i)
            out(i, j) = T(i, j);
    return out;
}

StlWrapper residual(StlWrapper const &A, StlWrapper const &B, CudaMultiVectorWrapper const &Xd,
                    HostDenseMatrix const &Td)
{
    StlWrapper X = to_host(Xd), T = to_host(Td);
    return A * X * T * X.transpose() + X * T * X.transpose() * A.transpose() + B * B.transpose();
}

struct ReferenceLoop
{
    typedef RAILS::Solver<CudaOperatorWrapper, CudaMultiVectorWrapper, HostDenseMatrix> type;
};
struct ProductLoop
{
    typedef RAILS::b200::Solver<CudaOperatorWrapper, CudaMultiVectorWrapper, HostDenseMatrix> type;
};

}

template<class Loop>
class SolverOnCudaWrappersTest : public testing::Test
{
};

typedef testing::Types<ReferenceLoop, ProductLoop> Loops;
TYPED_TEST_CASE(SolverOnCudaWrappersTest, Loops);

// :118-158 (StlSolver): random dense A, B = last entry of a random vector; solved twice
TYPED_TEST(SolverOnCudaWrappersTest, DenseSolver)
{
    int n = 20;
    StlWrapper A(n, n);
    A.random();
    StlWrapper B(n, 1);
    B.random();
    StlWrapper C = B.copy();
    C(n - 1, 0) = 0.0;
    B -= C;

    CudaOperatorWrapper Ad = to_device_operator(A);
    CudaMultiVectorWrapper Bd = to_device_multivector(B);
    CudaMultiVectorWrapper X(n, 1);
    HostDenseMatrix T;
    StlWrapper R_exp(n, n);
    R_exp = 0.0;

    typename TypeParam::type solver(Ad, Bd, Ad);
    int ret = solver.solve(X, T);
    EXPECT_EQ(0, ret);
    StlWrapper R = residual(A, B, X, T);
    EXPECT_VECTOR_NEAR(R_exp, R);

    ret = solver.solve(X, T);
    EXPECT_EQ(0, ret);
    R = residual(A, B, X, T);
    EXPECT_VECTOR_NEAR(R_exp, R);
}

// :202-232 (StlSolverRestart)
TYPED_TEST(SolverOnCudaWrappersTest, Restart)
{
    int n = 20;
    StlWrapper A, B;
    get_tridiagonal_problem(n, A, B);
    CudaOperatorWrapper Ad = to_device_operator(A);
    CudaMultiVectorWrapper Bd = to_device_multivector(B);
    CudaMultiVectorWrapper X(n, 1);
    HostDenseMatrix T;

    typename TypeParam::type solver(Ad, Bd, Ad);
    ParameterList params;
    params.set("Restart Size", 19);
    params.set("Reduced Size", 15);
    params.set("Expand Size", 1);
    params.set("Minimize solution space", false);
    solver.set_parameters(params);

    int ret = solver.solve(X, T);
    EXPECT_EQ(0, ret);
    StlWrapper R = residual(A, B, X, T);
    StlWrapper R_exp(n, n);
    R_exp = 0.0;
    EXPECT_GT(n, X.N());
    EXPECT_VECTOR_NEAR(R_exp, R);
}

// :234-275 (StlSolverMinimize)
TYPED_TEST(SolverOnCudaWrappersTest, Minimize)
{
    int n = 20;
    StlWrapper A, B;
    get_tridiagonal_problem(n, A, B);
    CudaOperatorWrapper Ad = to_device_operator(A);
    CudaMultiVectorWrapper Bd = to_device_multivector(B);
    CudaMultiVectorWrapper X(n, 1);
    HostDenseMatrix T;
    StlWrapper R_exp(n, n);
    R_exp = 0.0;

    typename TypeParam::type solver(Ad, Bd, Ad);
    ParameterList params;
    params.set("Minimize solution space", false);
    params.set("Tolerance", 1e-8);
    solver.set_parameters(params);

    int ret = solver.solve(X, T);
    EXPECT_EQ(0, ret);
    StlWrapper R = residual(A, B, X, T);
    EXPECT_EQ(n, X.N());
    EXPECT_VECTOR_NEAR(R_exp, R);

    params.set("Minimize solution space", true);
    solver.set_parameters(params);
    ret = solver.solve(X, T);
    EXPECT_EQ(0, ret);
    R = residual(A, B, X, T);
    EXPECT_GT(n, X.N());
    EXPECT_VECTOR_NEAR(R_exp, R);
}

// :277-310 (StlSolverRestartIterations)
TYPED_TEST(SolverOnCudaWrappersTest, RestartIterations)
{
    int n = 20;
    StlWrapper A, B;
    get_tridiagonal_problem(n, A, B);
    CudaOperatorWrapper Ad = to_device_operator(A);
    CudaMultiVectorWrapper Bd = to_device_multivector(B);
    CudaMultiVectorWrapper X(n, 1);
    HostDenseMatrix T;
    StlWrapper R_exp(n, n);
    R_exp = 0.0;

    typename TypeParam::type solver(Ad, Bd, Ad);
    ParameterList params;
    params.set("Restart iterations", 10);
    params.set("Minimize solution space", false);
    params.set("Expand size", 1);
    solver.set_parameters(params);

    int ret = solver.solve(X, T);
    EXPECT_EQ(0, ret);
    StlWrapper R = residual(A, B, X, T);
    EXPECT_GT(n, X.N());
    EXPECT_VECTOR_NEAR(R_exp, R);
}

// :312-352 (StlSolverRestartFromSolution): warm start after a change of A
TYPED_TEST(SolverOnCudaWrappersTest, RestartFromSolution)
{
    int n = 20;
    StlWrapper A, B;
    get_tridiagonal_problem(n, A, B);
    CudaOperatorWrapper Ad = to_device_operator(A);
    CudaMultiVectorWrapper Bd = to_device_multivector(B);
    CudaMultiVectorWrapper X(n, 1);
    HostDenseMatrix T;

    typename TypeParam::type solver(Ad, Bd, Ad);
    ParameterList params;
    params.set("Minimize solution space", true);
    params.set("Tolerance", 1e-8);
    solver.set_parameters(params);

    int ret = solver.solve(X, T);
    EXPECT_EQ(0, ret);
    EXPECT_GT(n, X.N());

    A(n - 1, n - 1) = 4.0;
    CudaOperatorWrapper Ad2 = to_device_operator(A);
    typename TypeParam::type solver2(Ad2, Bd, Ad2);
    params.set("Restart from solution", true);
    solver2.set_parameters(params);

    ret = solver2.solve(X, T);
    EXPECT_EQ(0, ret);
    StlWrapper R = residual(A, B, X, T);
    StlWrapper R_exp(n, n);
    R_exp = 0.0;
    EXPECT_GT(n, X.N());
    EXPECT_VECTOR_NEAR(R_exp, R);
}

#include "conformance_main.hpp"
