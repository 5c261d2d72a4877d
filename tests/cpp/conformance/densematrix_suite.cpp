// The reference's typed DenseMatrix suite, UNMODIFIED, with HostDenseMatrix added to
// its Types<> list (/root/reference/test/GenericDenseMatrixWrapper_test.cpp:52-58).
// Host only: runs in the build container as well.
#include "HostDenseMatrix.hpp"

#define RAILS_GTEST_EXTRA_TYPES , RAILS::HostDenseMatrix
#include "gtest/gtest.h"

#include "test/GenericDenseMatrixWrapper_test.cpp"

#define RAILS_CONFORMANCE_NO_GPU
#include "conformance_main.hpp"
