// The reference's typed MultiVector suite, UNMODIFIED, with the CUDA wrapper added
// to its Types<> list (the way the reference accepts a backend,
// /root/reference/test/GenericMultiVectorWrapper_test.cpp:52-58).  Built by
// oracle/build_ref.py into oracle/_ref/ (needs /root/reference), run on the GPU by
// tests/test_conformance.py.  StlWrapper runs alongside: it checks the gtest
// stand-in itself.
#include "CudaMultiVectorWrapper.hpp"

#define RAILS_GTEST_EXTRA_TYPES , RAILS::CudaMultiVectorWrapper
#include "gtest/gtest.h"

#include "test/GenericMultiVectorWrapper_test.cpp"

#include "conformance_main.hpp"
