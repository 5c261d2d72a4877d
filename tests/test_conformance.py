"""The reference's own test suites on the product's wrapper classes.

oracle/build_ref.py compiles, in the build container (where /root/reference
exists), the UNMODIFIED files
    /root/reference/test/GenericMultiVectorWrapper_test.cpp   (25 typed cases)
    /root/reference/test/GenericOperatorWrapper_test.cpp      (8 typed cases)
    /root/reference/test/GenericDenseMatrixWrapper_test.cpp   (10 typed cases)
against a single-header gtest stand-in (tests/cpp/gtest_shim) that appends
CudaMultiVectorWrapper / CudaOperatorWrapper / HostDenseMatrix to the suites'
`Types<...> Implementations` lists -- the way the reference accepts a backend --
plus a restatement of its algorithm tests (test/LyapunovSolver_test.cpp:118-352) for
both solver loops over the CUDA wrappers (tests/cpp/conformance/solver_suite.cpp).
The binaries travel to the GPU box in oracle/_ref/; StlWrapper runs in the same
binaries, which validates the stand-in itself.
"""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref")


def run_suite(name, expect_at_least):
    exe = os.path.join(REF, "conformance_" + name)
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/conformance_%s was not built (needs /root/reference at build time)" % name)
    r = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    out = r.stdout
    ran = re.search(r"\[==========\] (\d+) tests ran", out)
    passed = re.search(r"\[  PASSED  \] (\d+) tests", out)
    failed = re.findall(r"^\[  FAILED  \] (\S+)$", out, flags=re.M)
    assert ran and passed, out[-3000:]
    assert r.returncode == 0 and not failed, "failed: %s\n%s" % (sorted(set(failed)), out[-4000:])
    assert int(ran.group(1)) >= expect_at_least and int(passed.group(1)) == int(ran.group(1))
    return int(ran.group(1)), out


def test_dense_matrix_suite_on_host_dense_matrix():
    # 10 typed cases x {StlWrapper, HostDenseMatrix}; host only
    ran, out = run_suite("densematrix", 20)
    assert "GenericDenseMatrixWrapperTest/1.Eigs2" in out


@pytest.mark.gpu
def test_multivector_suite_on_cuda_multivector():
    # 25 typed cases x {StlWrapper, CudaMultiVectorWrapper}
    ran, out = run_suite("multivector", 50)
    assert "[       OK ] GenericMultiVectorWrapperTest/1.Orthogonalize" in out
    assert "[       OK ] GenericMultiVectorWrapperTest/1.View2" in out


@pytest.mark.gpu
def test_operator_suite_on_cuda_operator():
    # 8 typed cases x {StlWrapper, CudaOperatorWrapper + CudaMultiVectorWrapper + HostDenseMatrix}
    ran, out = run_suite("operator", 16)
    assert "[       OK ] GenericOperatorWrapperTest/1.Norm2" in out


@pytest.mark.gpu
def test_algorithm_tests_on_cuda_wrappers_both_loops():
    # 5 cases x {unmodified reference loop, product loop}
    ran, out = run_suite("solver", 10)
