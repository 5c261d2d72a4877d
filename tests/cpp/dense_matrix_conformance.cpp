// Driver of tests/test_host_sanitizers.py: HostDenseMatrix (rails_b200/host/
// HostDenseMatrix.cpp, the DenseMatrix concept of the B200 backend) against the
// known answers of the reference's typed conformance suite for that concept
// (/root/reference/test/GenericDenseMatrixWrapper_test.cpp:61-206 -- eigenvalues of
// [1 3; 3 4] at 1e-6, the same after a shrinking resize with stale data outside,
// fill, resize keeping entries and leading dimension, view assignment surviving
// a regrowth, infinity norm 7, transposed products at 1e-14, shapes under
// transpose), restated as plain checks and run under AddressSanitizer + UBSan.
#include "HostDenseMatrix.hpp"

#include <cmath>
#include <cstdio>

using RAILS::HostDenseMatrix;

static int failures = 0;
#define CHECK(cond)                                                                  \
    do {                                                                             \
        if (!(cond)) { std::printf("line %d: %s\n", __LINE__, #cond); ++failures; }  \
    } while (0)

static bool near(double a, double b, double tol) { return std::fabs(a - b) <= tol; }

int main()
{
    const double lo = (5.0 - std::sqrt(45.0)) / 2.0, hi = (5.0 + std::sqrt(45.0)) / 2.0;
    {   // eigenvalues of a symmetric 2 x 2 (:61-72)
        HostDenseMatrix A(2, 2), V(2, 2), D(2, 2);
        A(0, 0) = 1; A(0, 1) = 3; A(1, 0) = 3; A(1, 1) = 4;
        CHECK(A.eigs(V, D) == 0);
        CHECK(near(lo, D(0, 0), 1e-6) && near(hi, D(1, 0), 1e-6));
        // eigenvectors: A v = lambda v, orthonormal
        for (int e = 0; e < 2; ++e)
            for (int i = 0; i < 2; ++i)
                CHECK(near(A(i, 0) * V(0, e) + A(i, 1) * V(1, e), D(e, 0) * V(i, e), 1e-12));
    }
    {   // a 10 x 10 with entries outside the 4 x 4 it is shrunk to (:74-92)
        HostDenseMatrix D(10, 10), V(2, 2), W(2, 2);
        D = 0.0;
        D(0, 5) = 10.0; D(5, 0) = 10.0;
        D.resize(4, 4);
        D(0, 0) = 1; D(0, 1) = 3; D(1, 0) = 3; D(1, 1) = 4;
        CHECK(D.eigs(V, W) == 0);
        CHECK(W.M() == 4);
        double big = 0.0, small = 0.0;
        for (int i = 0; i < W.M(); ++i)
        {
            if (W(i, 0) > big) big = W(i, 0);
            if (W(i, 0) < small) small = W(i, 0);
        }
        CHECK(near(hi, big, 1e-6) && near(lo, small, 1e-6));
        // the `num` largest in magnitude above a tolerance (StlWrapper::eigs :433-479)
        HostDenseMatrix V2, W2;
        CHECK(D.eigs(V2, W2, 4, 1e-8) == 0);
        CHECK(W2.M() == 2 && near(std::fabs(W2(0, 0)), hi, 1e-6) && near(std::fabs(W2(1, 0)), -lo, 1e-6));
    }
    {   // fill (:94-103, :105-114)
        HostDenseMatrix A(2, 2), B(2, 2), D(10, 10);
        A = 2.0;
        for (int i = 0; i < 2; ++i)
            for (int j = 0; j < 2; ++j)
                CHECK(A(i, j) == 2.0);
        A = 0.0;
        D = 0.0; D(1, 1) = 1.0;
        D.resize(2, 2);
        D = 0.0;
        for (int i = 0; i < 2; ++i)
            for (int j = 0; j < 2; ++j)
                CHECK(D(i, j) == A(i, j));
    }
    {   // resize keeps entries; M, N, LDA follow (:116-127); an empty matrix can grow (:129-134)
        HostDenseMatrix A(2, 2);
        A = 0.0; A(0, 0) = 10.0; A(1, 1) = -3.0;
        A.resize(80, 100);
        CHECK(A.M() == 80 && A.N() == 100 && A.LDA() == 80 && A(0, 0) == 10.0 && A(1, 1) == -3.0);
        A(79, 99) = 5.0;
        A.resize(3, 3);
        A.resize(90, 120);
        CHECK(A(0, 0) == 10.0 && A(1, 1) == -3.0);
        HostDenseMatrix E;
        E.resize(10, 10);
        E(9, 9) = 1.0;
        CHECK(E.M() == 10 && E.N() == 10);
        // raw pointer + leading dimension, as handed to LAPACK (src/LyapunovSolver.hpp:357)
        double *raw = E;
        CHECK(raw[9 + 9 * E.LDA()] == 1.0);
    }
    {   // view assignment copies in place and survives a regrowth (:136-167)
        HostDenseMatrix A(2, 2), D(10, 10);
        A = 2.0; D = 1.0;
        D.resize(2, 2);
        D.view() = A;
        for (int i = 0; i < 2; ++i)
            for (int j = 0; j < 2; ++j)
                CHECK(D(i, j) == 2.0);
        D.resize(10, 10);
        for (int i = 0; i < 2; ++i)
            for (int j = 0; j < 2; ++j)
                CHECK(D(i, j) == 2.0);
        CHECK(D(5, 5) == 1.0);
        // copy() is deep, operator= of a non-view shares or copies but never aliases a temporary
        HostDenseMatrix C = A.copy();
        C(0, 0) = -1.0;
        CHECK(A(0, 0) == 2.0);
    }
    {   // norm_inf = 7 (:169-178); scaling
        HostDenseMatrix A(2, 2);
        A(0, 0) = 1; A(0, 1) = 3; A(1, 0) = -3; A(1, 1) = 4;
        CHECK(A.norm_inf() == 7.0);
        A *= 2.0;
        CHECK(A.norm_inf() == 14.0);
        A /= 4.0;
        CHECK(A(1, 1) == 2.0);
    }
    {   // transposed products (:180-197) and shapes (:199-205)
        HostDenseMatrix A(2, 2), a(2, 1);
        A(0, 0) = 1; A(0, 1) = 2; A(1, 0) = 3; A(1, 1) = 4;
        a(0, 0) = 1.2423; a(1, 0) = -4.9693;
        HostDenseMatrix b = A.transpose() * a, c = A * a;
        CHECK(near(a(0, 0) + 3.0 * a(1, 0), b(0, 0), 1e-14) && near(2.0 * a(0, 0) + 4.0 * a(1, 0), b(1, 0), 1e-14));
        CHECK(near(a(0, 0) + 2.0 * a(1, 0), c(0, 0), 1e-14) && near(3.0 * a(0, 0) + 4.0 * a(1, 0), c(1, 0), 1e-14));
        CHECK(a.M() == 2 && a.N() == 1 && a.transpose().M() == 1 && a.transpose().N() == 2);
        // X' VAV X of the restart compression (src/LyapunovSolver.hpp:283-297): 3 x 2 blocks
        HostDenseMatrix X(3, 2), S(3, 3);
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j)
                S(i, j) = 1.0 + i + 2 * j;
        for (int i = 0; i < 3; ++i) { X(i, 0) = i + 1.0; X(i, 1) = 1.0 - i; }
        HostDenseMatrix P = X.transpose() * (S * X);
        double want = 0.0;
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j)
                want += X(i, 1) * S(i, j) * X(j, 0);
        CHECK(P.M() == 2 && P.N() == 2 && near(P(1, 0), want, 1e-12));
    }
    std::printf(failures ? "FAILED (%d)\n" : "OK\n", failures);
    return failures != 0;
}
