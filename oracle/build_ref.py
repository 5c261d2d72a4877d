"""TEST INFRASTRUCTURE: build recipe for oracle/_ref (the reference arm).

Compiles the reference's own sources where they lie under /root/reference
(nothing is copied into the repo) together with the small C entry points in
oracle/ref_entry_*.cpp.  Outputs go only to oracle/_ref/ (git-ignored, but it
travels to the GPU box with the gpurun snapshot).

    oracle/_ref/librails_ref_stl.so   reference loop + reference StlWrapper
                                      (dense A; config C1 and the unit pins)
    oracle/_ref/librails_ref_csr.so   reference loop + reference StlWrapper +
                                      a CSR/OpenMP Operator (oracle/cpu_backend.cpp):
                                      the sparse CPU arm, any size, M = I

Flags follow the reference's no-build-type default `-O2` (CMakeLists.txt:16),
NOT its Release `-O3 -ffast-math` (CMakeLists.txt:18), so that the numbers are
reproducible.  SLICOT is absent on this image: sb03md_ is the
stand-in in rails_b200/host/sb03md_shim.cpp; BLAS/LAPACK is the OpenBLAS
bundled with opencv-python-headless.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = "/root/reference"
OUT = os.path.join(HERE, "_ref")


def openblas_path():
    import importlib.util
    spec = importlib.util.find_spec("cv2")
    cands = []
    if spec and spec.submodule_search_locations:
        site = os.path.dirname(list(spec.submodule_search_locations)[0])
        cands.append(os.path.join(site, "opencv_python_headless.libs"))
    cands.append("/opt/prime-rl/.venv/lib/python3.12/site-packages/"
                 "opencv_python_headless.libs")
    for d in cands:
        if os.path.isdir(d):
            for f in sorted(os.listdir(d)):
                if f.startswith("libopenblas"):
                    return d, os.path.join(d, f)
    raise RuntimeError("no LP64 OpenBLAS with plain LAPACK symbols found")


def run(cmd):
    print("+", " ".join(cmd), file=sys.stderr, flush=True)
    subprocess.check_call(cmd, stdout=sys.stderr)


def build(verbose=True):
    if not os.path.isdir(os.path.join(REF, "src")):
        if verbose:
            print("oracle/_ref: /root/reference absent, keeping prebuilt files")
        return False
    os.makedirs(OUT, exist_ok=True)
    libdir, blas = openblas_path()
    ref_srcs = [os.path.join(REF, "src", f) for f in
                ("StlWrapper.cpp", "StlVector.cpp", "LapackWrapper.cpp",
                 "SlicotWrapper.cpp", "Timer.cpp")]
    shim = os.path.join(ROOT, "rails_b200", "host", "sb03md_shim.cpp")
    common = ["g++", "-std=c++14", "-O2", "-fPIC", "-shared", "-fopenmp",
              "-I" + REF, "-I" + HERE, "-I" + os.path.join(ROOT, "rails_b200", "host")]
    link = [blas, "-Wl,--disable-new-dtags,-rpath," + libdir]

    run(common + ["-o", os.path.join(OUT, "librails_ref_stl.so"),
                  os.path.join(HERE, "ref_entry_stl.cpp")] + ref_srcs + [shim] + link)

    # reference loop over the product's CUDA wrappers (drop-in proof); links
    # the product libraries, which must have been built first
    prod = os.path.join(ROOT, "rails_b200")
    cuda_entry = os.path.join(HERE, "ref_entry_cuda.cpp")
    if os.path.exists(os.path.join(prod, "librails_host.so")):
        loop_srcs = [os.path.join(REF, "src", f) for f in
                     ("LapackWrapper.cpp", "SlicotWrapper.cpp", "Timer.cpp")]
        run(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-I" + REF, "-I" + HERE,
             "-I" + os.path.join(prod, "host"), "-I" + os.path.join(ROOT, "include"),
             "-o", os.path.join(OUT, "librails_refloop_cuda.so"), cuda_entry] + loop_srcs +
            ["-L" + prod, "-lrails_host", "-lrails_cuda", blas,
             "-Wl,--disable-new-dtags,-rpath," + libdir + ":$ORIGIN/../../rails_b200"])

    # The reference's own typed test suites, unmodified, with the product's wrapper
    # classes added to their Types<> lists through the gtest stand-in
    # (tests/cpp/gtest_shim), and its algorithm tests restated on those classes
    # (tests/cpp/conformance/).  Binaries: oracle/_ref/conformance_<suite>.
    conf = os.path.join(ROOT, "tests", "cpp", "conformance")
    if os.path.exists(os.path.join(prod, "librails_host.so")):
        inc = ["-I" + REF, "-I" + os.path.join(REF, "test"), "-I" + os.path.join(ROOT, "tests", "cpp", "gtest_shim"),
               "-I" + conf, "-I" + os.path.join(prod, "host"), "-I" + os.path.join(ROOT, "include")]
        stl = [os.path.join(REF, "src", f) for f in ("StlWrapper.cpp", "StlVector.cpp", "LapackWrapper.cpp",
                                                     "Timer.cpp")]
        for suite, extra in (("densematrix", []), ("multivector", []), ("operator", []),
                             ("solver", [os.path.join(REF, "src", "SlicotWrapper.cpp")])):
            run(["g++", "-std=c++17", "-O2"] + inc +
                ["-o", os.path.join(OUT, "conformance_" + suite), os.path.join(conf, suite + "_suite.cpp")] +
                stl + extra + ["-L" + prod, "-lrails_host", "-lrails_cuda", blas,
                               "-Wl,--disable-new-dtags,-rpath," + libdir + ":$ORIGIN/../../rails_b200"])

    # reference loop + reference StlWrapper (MultiVector, DenseMatrix) + CSR/OpenMP
    # Operator: the sparse CPU arm.  -ffp-contract=off keeps the operator's row sums
    # free of FMA contraction (the recipe's -O2 would contract on an FMA target).
    run(common + ["-ffp-contract=off", "-o", os.path.join(OUT, "librails_ref_csr.so"),
                  os.path.join(HERE, "ref_entry_csr.cpp"), os.path.join(HERE, "cpu_backend.cpp")] +
        ref_srcs + [shim] + link)
    return True


if __name__ == "__main__":
    build()
    sys.exit(0)
