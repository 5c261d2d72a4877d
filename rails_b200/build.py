"""Build recipe for the product libraries (in-tree, sm_100a only).

    rails_b200/librails_cuda.so   CUDA kernels + the C ABI of include/rails_cuda.h
    rails_b200/librails_host.so   C++ host side: wrapper classes, solver loop,
                                  sb03md stand-in, C entry points for Python
    rails_b200/rails_cuda_main    the command-line driver (A.mtx B.mtx M.mtx ->
                                  V.mtx T.mtx, counterpart of the reference's main.cpp)

nvcc cross-compiles without a GPU; the .so files travel to the GPU box with the
gpurun snapshot (they are git-ignored, not gpurun-ignored).
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
HOST = os.path.join(HERE, "host")
INC = os.path.join(ROOT, "include")
BUILD = os.path.join(HERE, "build")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3",
              "-std=c++17", "-Xcompiler", "-fPIC", "-I" + INC, "-I" + CSRC]


def openblas_path():
    d = os.path.join(os.path.dirname(os.path.dirname(os.__file__)), "site-packages")
    cands = []
    try:
        import importlib.util
        spec = importlib.util.find_spec("cv2")
        if spec and spec.submodule_search_locations:
            site = os.path.dirname(list(spec.submodule_search_locations)[0])
            cands.append(os.path.join(site, "opencv_python_headless.libs"))
    except Exception:
        pass
    cands.append("/opt/prime-rl/.venv/lib/python3.12/site-packages/opencv_python_headless.libs")
    for d in cands:
        if os.path.isdir(d):
            for f in sorted(os.listdir(d)):
                if f.startswith("libopenblas"):
                    return d, os.path.join(d, f)
    raise RuntimeError("no LP64 OpenBLAS with plain LAPACK symbols found")


def _run(cmd, quiet=False):
    if not quiet:
        print("+", " ".join(cmd), flush=True)
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        print(r.stdout)
        raise RuntimeError("command failed: " + " ".join(cmd))
    return r.stdout


def _stale(out, deps):
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    return any(os.path.getmtime(d) > t for d in deps)


def build_cuda(force=False, verbose=False):
    os.makedirs(BUILD, exist_ok=True)
    srcs = sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".h")]
    hdrs.append(os.path.join(INC, "rails_cuda.h"))
    objs, jobs = [], []
    for s in srcs:
        src = os.path.join(CSRC, s)
        obj = os.path.join(BUILD, s[:-3] + ".o")
        objs.append(obj)
        if force or _stale(obj, [src] + hdrs):
            cmd = ["nvcc"] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
            jobs.append(cmd)
    with ThreadPoolExecutor(max_workers=8) as ex:
        for out in ex.map(_run, jobs):
            if verbose:
                print(out)
    lib = os.path.join(HERE, "librails_cuda.so")
    if force or jobs or _stale(lib, objs):
        _run(["nvcc", "-shared", "-o", lib] + objs + ["-lnccl", "-lcudart"])
    return lib


def build_host(force=False):
    libdir, blas = openblas_path()
    srcs = sorted(os.path.join(HOST, f) for f in os.listdir(HOST) if f.endswith(".cpp"))
    hdrs = [os.path.join(HOST, f) for f in os.listdir(HOST) if f.endswith((".hpp", ".h"))]
    hdrs.append(os.path.join(INC, "rails_cuda.h"))
    lib = os.path.join(HERE, "librails_host.so")
    if force or _stale(lib, srcs + hdrs + [os.path.join(HERE, "librails_cuda.so")]):
        _run(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-I" + INC, "-I" + HOST, "-o", lib]
             + srcs + ["-L" + HERE, "-lrails_cuda", blas,
                       "-Wl,--disable-new-dtags,-rpath," + libdir + ":$ORIGIN"])
    return lib


def build_cli(force=False):
    libdir, blas = openblas_path()
    src = os.path.join(HERE, "cli", "rails_main.cpp")
    hdrs = [os.path.join(HOST, f) for f in os.listdir(HOST) if f.endswith((".hpp", ".h"))]
    exe = os.path.join(HERE, "rails_cuda_main")
    if force or _stale(exe, [src] + hdrs + [os.path.join(HERE, "librails_host.so")]):
        _run(["g++", "-std=c++17", "-O2", "-I" + INC, "-I" + HOST, "-o", exe, src,
              "-L" + HERE, "-lrails_host", "-lrails_cuda", blas, "-lpthread",
              "-Wl,--disable-new-dtags,-rpath," + libdir + ":$ORIGIN"])
    return exe


def build_all(force=False, verbose=False):
    build_cuda(force, verbose)
    build_host(force)
    build_cli(force)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
