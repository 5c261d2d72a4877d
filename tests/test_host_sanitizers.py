"""Host code under AddressSanitizer + UndefinedBehaviorSanitizer (SURVEY.md §5:
"ASan/UBSan on the host shim"): the sparse LU factorisation and its orderings,
the newest host code with the most pointer arithmetic (depth-first reach, merges
of adjacency lists, the dissection stack)."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(shutil.which("g++") is None, reason="no g++")
def test_sparse_lu_is_clean_under_asan_and_ubsan(tmp_path):
    exe = str(tmp_path / "sparse_lu_sanitize")
    cmd = ["g++", "-std=c++17", "-O1", "-g", "-fsanitize=address,undefined", "-fno-sanitize-recover=all",
           "-fno-omit-frame-pointer", "-I" + os.path.join(ROOT, "rails_b200", "host"),
           os.path.join(ROOT, "tests", "cpp", "sparse_lu_sanitize.cpp"),
           os.path.join(ROOT, "rails_b200", "host", "SparseLU.cpp"), "-o", exe]
    b = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if b.returncode != 0 and "sanitize" in b.stdout:
        pytest.skip("sanitizer runtime not available: " + b.stdout[-200:])
    assert b.returncode == 0, b.stdout[-2000:]
    r = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:]
    assert r.stdout.strip().endswith("OK")
    assert "runtime error" not in r.stdout and "AddressSanitizer" not in r.stdout


@pytest.mark.skipif(shutil.which("nvcc") is None, reason="no nvcc")
def test_device_schedule_planner_is_clean_under_asan(tmp_path):
    """plan_triangle (host code inside csrc/sptrsv.cu) through rails_lu_plan_emulate:
    no CUDA call, so it runs here; compiled with the host compiler's ASan."""
    exe = str(tmp_path / "lu_plan_sanitize")
    # sptrsv.cu and what it links against (context, reductions, blas-1)
    srcs = [os.path.join(ROOT, "rails_b200", "csrc", f)
            for f in ("sptrsv.cu", "context.cu", "exchange.cu", "blas1.cu")]
    cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O1", "-g", "-std=c++17", "-w",
           "-Xcompiler", "-fsanitize=address,-fno-omit-frame-pointer,-fPIC",
           "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(ROOT, "rails_b200", "csrc"),
           "-I" + os.path.join(ROOT, "rails_b200", "host"),
           os.path.join(ROOT, "tests", "cpp", "lu_plan_sanitize.cpp"),
           os.path.join(ROOT, "rails_b200", "host", "SparseLU.cpp")] + srcs + ["-lnccl", "-o", exe]
    b = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if b.returncode != 0 and ("sanitize" in b.stdout or "asan" in b.stdout):
        pytest.skip("sanitizer runtime not available: " + b.stdout[-200:])
    assert b.returncode == 0, b.stdout[-2000:]
    env = dict(os.environ, ASAN_OPTIONS="protect_shadow_gap=0:detect_leaks=0")
    r = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-3000:]
    assert r.stdout.strip().endswith("OK") and "AddressSanitizer" not in r.stdout


@pytest.mark.skipif(shutil.which("g++") is None, reason="no g++")
def test_file_readers_survive_corrupt_input_under_asan_and_ubsan(tmp_path):
    """MatrixMarket and parameter-file readers on valid, truncated, bit-flipped
    and size-lying files: an error code or a consistent result, no exception, no
    out-of-bounds access.  (Found on its first run: a size line promising 2e9
    entries made read_crs reserve 48 GB and die of std::length_error.)"""
    exe = str(tmp_path / "parsers_sanitize")
    host = os.path.join(ROOT, "rails_b200", "host")
    cmd = ["g++", "-std=c++17", "-O1", "-g", "-fsanitize=address,undefined", "-fno-sanitize-recover=all",
           "-fno-omit-frame-pointer", "-I" + host, os.path.join(ROOT, "tests", "cpp", "parsers_sanitize.cpp"),
           os.path.join(host, "MatrixMarket.cpp"), os.path.join(host, "ParameterFile.cpp"), "-o", exe]
    b = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if b.returncode != 0 and "sanitize" in b.stdout:
        pytest.skip("sanitizer runtime not available: " + b.stdout[-200:])
    assert b.returncode == 0, b.stdout[-2000:]
    work = tmp_path / "files"
    work.mkdir()
    r = subprocess.run([exe, str(work)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:]
    assert r.stdout.strip().endswith("OK")
    assert "runtime error" not in r.stdout and "AddressSanitizer" not in r.stdout


REF = "/root/reference"


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "src")), reason="reference tree absent (GPU box)")
def test_solver_loop_is_clean_under_asan(tmp_path):
    """The product's solver loop (RailsSolverImpl.hpp, RailsCoefficientSpace.hpp: the
    MATLAB options, residual methods 1 and 2, projection methods) over the
    reference's StlWrapper, built with AddressSanitizer; a subset of the CPU solver
    tests is run against that build in a child process."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import build_ref
    libdir, blas = build_ref.openblas_path()
    lib = str(tmp_path / "librails_ref_stl_asan.so")
    srcs = [os.path.join(REF, "src", f) for f in
            ("StlWrapper.cpp", "StlVector.cpp", "LapackWrapper.cpp", "SlicotWrapper.cpp", "Timer.cpp")]
    cmd = ["g++", "-std=c++14", "-O2", "-g", "-fPIC", "-shared", "-fopenmp", "-fsanitize=address",
           "-fno-omit-frame-pointer", "-I" + REF, "-I" + os.path.join(ROOT, "oracle"),
           "-I" + os.path.join(ROOT, "rails_b200", "host"), "-o", lib,
           os.path.join(ROOT, "oracle", "ref_entry_stl.cpp")] + srcs + [
           os.path.join(ROOT, "rails_b200", "host", "sb03md_shim.cpp"), blas,
           "-Wl,--disable-new-dtags,-rpath," + libdir]
    b = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if b.returncode != 0 and "sanitize" in b.stdout:
        pytest.skip("sanitizer runtime not available: " + b.stdout[-200:])
    assert b.returncode == 0, b.stdout[-2000:]
    asan = subprocess.run(["g++", "-print-file-name=libasan.so"], stdout=subprocess.PIPE, text=True).stdout.strip()
    if not os.path.isabs(asan) or not os.path.exists(asan):
        pytest.skip("libasan.so not found")
    env = dict(os.environ, LD_PRELOAD=asan, ASAN_OPTIONS="detect_leaks=0:verify_asan_link_order=0",
               RAILS_REF_STL_LIB=lib, OPENBLAS_NUM_THREADS="1")
    tests = ["tests/test_solver_loop.py",
             "tests/test_schur_oracle.py::test_projection_methods_follow_matlab_restatement",
             "tests/test_matlab_features.py::test_product_loop_follows_matlab_restatement",
             "tests/test_matlab_features.py::test_exact_residual_eigenpairs_at_every_iteration"]
    r = subprocess.run([sys.executable, "-m", "pytest", "-q", "-x", "-p", "no:cacheprovider"] + tests,
                       cwd=ROOT, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1200)
    tail = "\n".join(line for line in r.stdout.splitlines() if not line.startswith("Warning"))[-3000:]
    assert "AddressSanitizer" not in r.stdout, tail
    assert r.returncode == 0, tail


@pytest.mark.skipif(shutil.which("g++") is None, reason="no g++")
def test_host_dense_matrix_conformance_under_asan_and_ubsan(tmp_path):
    """HostDenseMatrix against the known answers of the reference's typed conformance
    suite for the DenseMatrix concept (test/GenericDenseMatrixWrapper_test.cpp:61-206),
    under the sanitizers."""
    from rails_b200 import build as product_build
    libdir, blas = product_build.openblas_path()
    exe = str(tmp_path / "dense_matrix_conformance")
    host = os.path.join(ROOT, "rails_b200", "host")
    cmd = ["g++", "-std=c++17", "-O1", "-g", "-fsanitize=address,undefined", "-fno-sanitize-recover=all",
           "-fno-omit-frame-pointer", "-I" + host, os.path.join(ROOT, "tests", "cpp", "dense_matrix_conformance.cpp"),
           os.path.join(host, "HostDenseMatrix.cpp"), blas, "-Wl,--disable-new-dtags,-rpath," + libdir, "-o", exe]
    b = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if b.returncode != 0 and "sanitize" in b.stdout:
        pytest.skip("sanitizer runtime not available: " + b.stdout[-200:])
    assert b.returncode == 0, b.stdout[-2000:]
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=0", OPENBLAS_NUM_THREADS="1",
               LD_LIBRARY_PATH=libdir + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
    r = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stdout[-3000:]
    assert r.stdout.strip().endswith("OK")
    assert "runtime error" not in r.stdout and "AddressSanitizer" not in r.stdout
