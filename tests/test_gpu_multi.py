"""Row-partitioned path on >= 2 GPUs (skipped on a single-GPU box): one rank
per GPU under torch.distributed.run; see tests/dist_worker.py for the checks."""
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

pytestmark = pytest.mark.gpu


def _gpu_count():
    try:
        out = subprocess.run(["nvidia-smi", "-L"], stdout=subprocess.PIPE, text=True, timeout=30).stdout
        return sum(1 for l in out.splitlines() if l.startswith("GPU "))
    except Exception:
        return 0


@pytest.mark.parametrize("world", [2, 4])
def test_row_partitioned_solve(world):
    if _gpu_count() < world:
        pytest.skip("needs %d GPUs" % world)
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                        "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
                        "--master-port", str(port), os.path.join(ROOT, "tests", "dist_worker.py")],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0 and "DIST_OK" in r.stdout, r.stdout[-4000:]


def test_cli_two_ranks(tmp_path):
    """rails_cuda_main started once per GPU with RANK / WORLD_SIZE / LOCAL_RANK:
    the ranks meet through the rendezvous file, each reads its row block of the
    .mtx files, rank 0 writes one V.mtx.  Checked on the explicit residual and
    against the single-rank run of the same files."""
    if _gpu_count() < 2:
        pytest.skip("needs 2 GPUs")
    import numpy as np
    import scipy.sparse as sp
    sys.path.insert(0, ROOT)
    from rails_b200 import api, problems
    import harness
    d = str(tmp_path)
    g, n_b = 24, 2
    m = g * g
    rp, ci, v = problems.laplacian(g, 2)
    api.mm_write_crs(os.path.join(d, "A.mtx"), rp, ci, v, m)
    B = np.random.RandomState(7).uniform(-1, 1, (m, n_b))
    Bs = sp.csr_matrix(B)
    api.mm_write_crs(os.path.join(d, "B.mtx"), Bs.indptr.astype(np.int32), Bs.indices.astype(np.int32),
                     Bs.data, n_b)
    open(os.path.join(d, "p.ini"), "w").write(
        "[Lyapunov Solver]\nTolerance = 1e-6\nExpand size = 2\nLanczos iterations = 12\n"
        "[Eigenvalue Solver]\nNumber of Eigenvalues = 4\n")
    exe = os.path.join(ROOT, "rails_b200", "rails_cuda_main")
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", LOCAL_RANK=str(r), MASTER_PORT=str(port))
        procs.append(subprocess.Popen([exe, os.path.join(d, "p.ini"), "--dir", d], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=600)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs[0][-3000:] + outs[1][-3000:]
    assert "Profile" in outs[0] and "Profile" not in outs[1]
    V = api.mm_read_dense(os.path.join(d, "V.mtx"))
    T = api.mm_read_dense(os.path.join(d, "T.mtx"))
    assert V.shape[0] == m and not [f for f in os.listdir(d) if ".part" in f]
    A = harness.csr_to_dense(rp, ci, v, m)
    X = V @ T @ V.T
    R = A @ X + X @ A.T + B @ B.T
    assert np.linalg.norm(R, 2) < 1e-5 * np.linalg.norm(B.T @ B, 2)
    r1 = subprocess.run([exe, os.path.join(d, "p.ini"), "--dir", d, "--V", "V1.mtx", "--T", "T1.mtx"],
                        stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r1.returncode == 0, r1.stdout[-3000:]
    V1 = api.mm_read_dense(os.path.join(d, "V1.mtx"))
    T1 = api.mm_read_dense(os.path.join(d, "T1.mtx"))
    assert harness.lowrank_rel_diff(V, T, V1, T1) < 1e-3
