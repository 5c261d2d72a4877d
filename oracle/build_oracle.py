"""TEST INFRASTRUCTURE: builds oracle/liboracle.so, the CPU restatement.

Flags: -O2 -ffp-contract=off (the reference's default build has no FMA
contraction), OpenMP for the row loops, the image's LP64 OpenBLAS
for the k x k LAPACK calls (dsyev, dgees, dtrsyl).
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from build_ref import openblas_path  # noqa: E402

SRCS = ["oracle_kernels.cpp", "oracle_solver.cpp", "oracle_solve.cpp"]


def build(force=False):
    out = os.path.join(HERE, "liboracle.so")
    srcs = [os.path.join(HERE, s) for s in SRCS]
    if not force and os.path.exists(out) and all(
            os.path.getmtime(s) <= os.path.getmtime(out) for s in srcs):
        return out
    libdir, blas = openblas_path()
    cmd = ["g++", "-std=c++14", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-fopenmp",
           "-o", out] + srcs + [blas, "-Wl,--disable-new-dtags,-rpath," + libdir]
    print("+", " ".join(cmd), file=sys.stderr, flush=True)
    subprocess.check_call(cmd, stdout=sys.stderr)
    return out


if __name__ == "__main__":
    build(force=True)
