"""One launch each of the kernels VERDICT r01 asked ncu evidence for, at the C3 row
count: SpMM with e = 4 and e = 16 right-hand sides, and the CGS2 chain of one column
against k = 32 (rails_mv_orthogonalize).  Run under
    ncu --set full -k regex:'spmm_kernel|dot_kernel|gemm_kernel' ...
"""
import ctypes, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rails_b200 import api, problems
api.init(0)
lib = api.cuda()
g = int(sys.argv[1]) if len(sys.argv) > 1 else 256
m = g ** 3
rp, ci, v = problems.build("laplacian", g, 3)
A = ctypes.c_void_p()
api.check(lib.rails_csr_create(m, m, m, 0, api.iptr(rp), api.iptr(ci), api.dptr(v), ctypes.byref(A)), "csr_create")
k = 32
V = api.MultiVector(m, k + 1)
X = api.MultiVector(m, 16)
Y = api.MultiVector(m, 16)
api.check(lib.rails_mv_random(V.h, 0, k + 1, 0, ctypes.c_uint(7)), "random")
api.check(lib.rails_mv_random(X.h, 0, 16, 0, ctypes.c_uint(8)), "random")
for e in (4, 16):
    api.check(lib.rails_csr_spmm(A, 0, X.h, 0, e, Y.h, 0), "spmm")
api.check(lib.rails_mv_orthogonalize(V.h, 0, k, k), "cgs2")
lib.rails_cuda_sync()
print("done", m, int(rp[-1]))
