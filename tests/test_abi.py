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


def test_lanczos_step_launch_shapes_fit_the_machine():
    """Host-only plan of the one-kernel Lanczos step (rails_lanczos_step_shape, no CUDA
    call): for every space size and block shape the solver can present, the chosen tile,
    pipeline depth and warp roles respect the limits the kernel is written for -- 216 KB
    of dynamic shared memory, at most 16 consumer warps (8 for the wide instances), 2-6
    stages, tiles of whole 128-row TMA boxes, at most 64 TMA requests per slab -- and the
    path is declined (tile rows 0) beyond 96 block columns, where the two-kernel chain
    runs."""
    lib = api.cuda()
    shape = (ctypes.c_int * 6)()
    seen = set()
    declined = []
    for n_b in (1, 2, 3, 4, 16):
        for k in range(0, 60):
            for blocks in ((k, k, n_b), (k, k, n_b, 0)):          # [V AV B]; an empty fourth block
                if sum(blocks) == 0:
                    continue
                xn = (ctypes.c_int * len(blocks))(*blocks)
                assert lib.rails_lanczos_step_shape(len(blocks), xn, shape) == 0
                tr, st, groups, smem, nreq, warps = list(shape)
                total = 2 * k + n_b
                if total > 96:
                    assert tr == 0 and smem == 0
                    continue
                if tr == 0:
                    # no candidate fits the shared-memory budget: only right at the column
                    # limit with padded blocks (e.g. 46 + 46 + 4); the two-kernel chain runs
                    assert total >= 90, (k, n_b)
                    declined.append((k, n_b))
                    continue
                assert tr in (128, 256, 384, 512), (k, n_b, tr)
                assert 2 <= st <= 6
                assert smem <= 216 * 1024
                assert nreq <= 64
                assert groups == (1 if total <= 50 else -(-total // 24))
                assert warps == tr // 32 * groups
                assert warps <= (8 if 24 < total <= 50 else 16)
                # the staged tiles alone must fit: stages x (padded columns + 2 basis columns) x rows
                assert st * (total + 2) * tr * 8 <= smem
                seen.add((tr, st, groups))
    # the shapes the measurements of profiles/r02_lanczos_step.md were taken at
    for k, want in ((5, (512, 3, 1)), (15, (256, 3, 1)), (24, (256, 2, 1)), (30, (128, 3, 3)), (47, (128, 2, 4))):
        xn = (ctypes.c_int * 3)(k, k, 1)
        assert lib.rails_lanczos_step_shape(3, xn, shape) == 0
        assert tuple(shape[:3]) == want, (k, tuple(shape[:3]))
    assert len(seen) >= 5
    assert len(set(declined)) <= 4, declined
    assert lib.rails_lanczos_step_shape(0, None, shape) == 2          # RAILS_ERR_ARG
