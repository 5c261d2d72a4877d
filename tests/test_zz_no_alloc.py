"""rails_mv_dot_dev_noalloc: the allocation-free X'Y (static device scratch, no
counterpart in the reference)
against the existing rails_mv_dot_dev, through the C ABI.

Collected last (test_zz_*): a failure here must
not hide the parity tests.  Self-contained: it passes alone in a fresh process and
after the other tests (it shares nothing with them but the initialised library).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _both(gpu, X, Y):
    """(X'Y by rails_mv_dot_dev, X'Y by rails_mv_dot_dev_noalloc), as numpy a x b."""
    lib = gpu.cuda()
    a, b = X.shape[1], Y.shape[1]
    Xd, Yd = gpu.MultiVector.from_numpy(X), gpu.MultiVector.from_numpy(Y)
    ref = gpu.MultiVector(a, b, replicated=True)
    new = gpu.MultiVector(a, b + 1, replicated=True)      # written at column offset 1
    gpu.check(lib.rails_mv_fill(new.h, 0, b + 1, -7.0), "fill")
    gpu.check(lib.rails_mv_dot_dev(Xd.h, 0, a, Yd.h, 0, b, ref.h, 0), "dot_dev")
    in_use = lib.rails_cuda_bytes_in_use()
    gpu.check(lib.rails_mv_dot_dev_noalloc(Xd.h, 0, a, Yd.h, 0, b, new.h, 1), "dot_dev_noalloc")
    assert lib.rails_cuda_bytes_in_use() == in_use
    got = new.numpy()
    assert np.all(got[:, 0] == -7.0)                      # nothing written outside the block
    out = ref.numpy(), got[:, 1:]
    for mv in (Xd, Yd, ref, new):
        mv.free()
    return out


def test_noalloc_dot_matches_dot_dev_on_two_calls_in_a_row(gpu):
    rs = np.random.RandomState(11)
    # first call: k x 1 (the shape of the Lanczos / CGS2 projections), ragged row count
    m = 300001
    X, Y = rs.uniform(-1, 1, (m, 37)), rs.uniform(-1, 1, (m, 1))
    ref, got = _both(gpu, X, Y)
    exact = X.T @ Y
    assert np.max(np.abs(ref - exact)) < 1e-12 * m       # the bar tests/test_gpu_kernels.py uses
    assert np.max(np.abs(got - exact)) < 1e-12 * m
    assert np.max(np.abs(got - ref)) <= 1e-13 * np.max(np.abs(ref))
    # second call straight after it, different data, SMALLER shape: the array still
    # holds the first call's partials, which must not leak into this result
    m2 = 4099
    X2, Y2 = rs.uniform(-1, 1, (m2, 3)), rs.uniform(-1, 1, (m2, 6))    # held block on the X side (swap)
    ref2, got2 = _both(gpu, X2, Y2)
    exact2 = X2.T @ Y2
    assert np.max(np.abs(got2 - exact2)) < 1e-12 * m2
    assert np.max(np.abs(got2 - ref2)) <= 1e-13 * np.max(np.abs(ref2))
    # and the first problem again: same answer as before
    ref3, got3 = _both(gpu, X, Y)
    assert np.array_equal(got3, got)


def test_noalloc_dot_rejects_what_exceeds_the_static_scratch(gpu):
    lib = gpu.cuda()
    # 592 CTAs x 40 streamed x 8 held columns = 189 440 doubles > 151 552
    m = 310000
    Xd, Yd = gpu.MultiVector(m, 40), gpu.MultiVector(m, 8)
    out = gpu.MultiVector(40, 8, replicated=True)
    errors = lib.rails_cuda_error_count()
    assert lib.rails_mv_dot_dev_noalloc(Xd.h, 0, 40, Yd.h, 0, 8, out.h, 0) == 2      # RAILS_ERR_ARG
    assert b"static array holds" in lib.rails_cuda_last_error()
    assert lib.rails_cuda_error_count() == errors + 1
    # the same operands, one held column at a time, fit
    assert lib.rails_mv_dot_dev_noalloc(Xd.h, 0, 40, Yd.h, 0, 1, out.h, 0) == 0
    # shape errors are reported like the existing entry point's
    assert lib.rails_mv_dot_dev_noalloc(Xd.h, 0, 40, Yd.h, 0, 1, out.h, 8) == 2
    for mv in (Xd, Yd, out):
        mv.free()
