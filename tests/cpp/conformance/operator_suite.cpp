// The reference's typed Operator suite, UNMODIFIED, with the CUDA wrappers added to
// its Types<> list (/root/reference/test/GenericOperatorWrapper_test.cpp:64-69).
#include "CudaMultiVectorWrapper.hpp"
#include "CudaOperatorWrapper.hpp"
#include "HostDenseMatrix.hpp"

// declared before the stand-in is included so that the extra list entry can name it;
// the test file defines it
template <typename A, typename B, typename C>
struct TypeDefinitions;

#define RAILS_GTEST_EXTRA_TYPES \
    , TypeDefinitions<RAILS::CudaOperatorWrapper, RAILS::CudaMultiVectorWrapper, RAILS::HostDenseMatrix>
#include "gtest/gtest.h"

#include "test/GenericOperatorWrapper_test.cpp"

#include "conformance_main.hpp"
