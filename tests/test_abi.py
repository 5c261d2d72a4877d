"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports
every symbol include/rails_cuda.h declares, and refuses to compute without a
GPU (no CPU fallback)."""
import ctypes
import os
import re

import pytest

from rails_b200 import api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "rails_cuda.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rails_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported_and_bound():
    lib = api.cuda()
    names = declared_symbols()
    assert len(names) >= 40
    bound = {n for n, _, _ in api.CUDA_SYMBOLS}
    for n in names:
        assert hasattr(lib, n), "librails_cuda.so does not export " + n
        assert n in bound, "rails_b200/api.py does not bind " + n
    assert bound <= set(names), "api.py binds symbols the header does not declare"


def test_host_library_loads():
    lib = api.host()
    assert hasattr(lib, "rails_host_solve")
    assert hasattr(lib, "rails_host_last_sections")


def test_product_does_not_link_the_oracle():
    import subprocess
    for so in ("librails_cuda.so", "librails_host.so"):
        out = subprocess.run(["ldd", os.path.join(ROOT, "rails_b200", so)],
                             stdout=subprocess.PIPE, text=True).stdout
        assert "liboracle" not in out and "librails_ref" not in out
    for dirpath, _, files in os.walk(os.path.join(ROOT, "rails_b200")):
        for f in files:
            if f.endswith((".py", ".cpp", ".hpp", ".h", ".cu")):
                src = open(os.path.join(dirpath, f)).read()
                assert "liboracle" not in src and "import harness" not in src, f


def _no_gpu():
    try:
        import torch
        return not torch.cuda.is_available()
    except Exception:
        return True


@pytest.mark.skipif(not _no_gpu(), reason="a GPU is present")
def test_fails_loudly_without_gpu():
    lib = api.cuda()
    assert lib.rails_cuda_init(0) == 1                      # RAILS_ERR_CUDA
    assert b"no CPU fallback" in lib.rails_cuda_last_error()
    h = ctypes.c_void_p()
    assert lib.rails_mv_alloc(10, 1, 0, ctypes.byref(h)) == 3   # RAILS_ERR_STATE
    with pytest.raises(api.RailsError):
        api.init(0)
    # the singular-mass path as well: no device solves, no Schur operator, no solve
    import numpy as np
    one_i, one_d = np.zeros(2, dtype=np.int32), np.ones(1)
    assert lib.rails_lu_create(1, api.iptr(one_i), api.iptr(one_i), api.dptr(one_d), api.iptr(np.array([0, 1], dtype=np.int32)),
                               api.iptr(one_i), api.dptr(one_d), api.iptr(one_i), api.iptr(one_i), ctypes.byref(h)) == 3
    assert lib.rails_idx_create(1, api.iptr(one_i), ctypes.byref(h)) == 3
    t = ctypes.c_double(0)
    assert lib.rails_cuda_fp64_peak(0, ctypes.byref(t)) == 3
    rp, ci, v = np.array([0, 1, 2], dtype=np.int32), np.array([0, 1], dtype=np.int32), np.array([-1.0, -2.0])
    with pytest.raises(api.RailsError):
        api.SchurOperator(rp, ci, v, np.array([0.0, 1.0]))
    q = api.solve_schur(rp, ci, v, np.array([0.0, 1.0]), np.ones((2, 1)), {"Tolerance": 1e-4})
    assert q["ret"] == 200


def test_partition_rows_matches_epetra_linear_map():
    lib = api.cuda()
    from rails_b200 import problems
    LL = ctypes.c_longlong
    for m, P in [(10, 3), (16777216, 8), (7, 8), (0, 2), (134217728, 8)]:
        prev = 0
        for r in range(P):
            b, e = LL(0), LL(0)
            assert lib.rails_partition_rows(m, P, r, ctypes.byref(b), ctypes.byref(e)) == 0
            assert (b.value, e.value) == problems.partition_rows(m, P, r)
            assert b.value == prev
            prev = e.value
            assert e.value - b.value in (m // P, m // P + 1)
        assert prev == m
