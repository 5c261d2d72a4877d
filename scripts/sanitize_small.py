"""Small instances of this round's new kernels for compute-sanitizer (memcheck,
racecheck, synccheck): the one-kernel Lanczos step in every instance (one column group,
wide, several groups), the fused CGS2, the device halo plan, the device transpose and a
short solve.  Sizes are chosen with ragged tails (rows not a multiple of the tile)."""
import ctypes, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rails_b200 import api, problems
api.init(0); api.host()
lib = api.cuda()
P = ctypes.c_void_p
rows = 20011
for k in (3, 10, 15, 22, 30, 45):
    n, L = 1, 5
    V, AV, B, Q = (api.MultiVector(rows, k), api.MultiVector(rows, k), api.MultiVector(rows, n),
                   api.MultiVector(rows, L + 1))
    for mv, c, s in ((V, k, 3), (AV, k, 4), (B, n, 5), (Q, 1, 9)):
        api.check(lib.rails_mv_random(mv.h, 0, c, 0, ctypes.c_uint(s)), "random")
        api.check(lib.rails_mv_scal(mv.h, 0, c, 1.0 / np.sqrt(rows)), "scal")
    T = np.asfortranarray(-1e-2 * np.eye(k))
    hs = (P * 3)(V.h, AV.h, B.h)
    c0 = (ctypes.c_int * 3)(0, 0, 0)
    nn = (ctypes.c_int * 3)(k, k, n)
    al, be = np.zeros(L), np.zeros(L)
    steps = ctypes.c_int(0)
    api.check(lib.rails_mv_lanczos_run(3, hs, c0, nn, api.dptr(T), k, Q.h, 0, L, api.dptr(al), api.dptr(be),
                                       ctypes.byref(steps)), "lanczos_run")
    # CGS2 of the last column of V against the others (fused path)
    api.check(lib.rails_mv_orthogonalize(V.h, 0, k - 1, k - 1), "cgs2")
    for mv in (V, AV, B, Q):
        mv.free()
    print("k", k, "steps", steps.value, flush=True)
# device halo plan + device transpose on a small grid block
g = 14
m = g ** 3
r0, r1 = problems.partition_rows(m, 3, 1)
rp, ci, v = problems.build("convdiff", g, 3, r0, r1)
nnz = int(rp[-1])
lcol = np.zeros(nnz, dtype=np.int32); gg = np.zeros(nnz, dtype=np.int32); gl = np.zeros(nnz, dtype=np.int32)
need = np.zeros(3, dtype=np.int32); ng = ctypes.c_int(0)
api.check(lib.rails_csr_halo_plan_device(r1 - r0, m, r0, 3, api.iptr(rp), api.iptr(ci), api.iptr(lcol), ctypes.byref(ng),
                                         api.iptr(gg), api.iptr(gl), api.iptr(need)), "plan")
rpf, cif, vf = problems.build("convdiff", g, 3)
A = P()
api.check(lib.rails_csr_create(m, m, m, 0, api.iptr(rpf), api.iptr(cif), api.dptr(vf), ctypes.byref(A)), "csr")
X, Y = api.MultiVector(m, 2), api.MultiVector(m, 2)
api.check(lib.rails_mv_random(X.h, 0, 2, 0, ctypes.c_uint(1)), "random")
api.check(lib.rails_csr_spmm(A, 1, X.h, 0, 2, Y.h, 0), "spmm^T")
lib.rails_csr_free(A)
# a short solve through the product loop (chain mode)
o = api.solve(rpf, cif, vf, 1, {"Tolerance": 1e-4, "Expand size": 1, "Lanczos iterations": 6, "Maximum iterations": 12})
print("solve", o["ret"], o["iterations"], flush=True)
lib.rails_cuda_sync()
print("done")
