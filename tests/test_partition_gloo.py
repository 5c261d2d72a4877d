"""CPU tests (gloo, world_size 2 and 3) of the host side of the row-partitioned
path: the Epetra-style linear map, the halo plan rails_csr_halo_plan builds for
a rank's rows (column renumbering [owned | ghost], ghost lists grouped by
owner), and the id exchange rails_csr_create performs over NCCL -- replayed here
with gloo send/recv pairs.  The distributed SpMM assembled from the plan must equal
the global SpMM bit for bit (same CSR order within every row); the transposed
apply (reverse halo: ghost contributions sent back to the owners and added) must
equal the global A'X.  No GPU and no compute entry point of the library is used."""
import ctypes
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _csr_apply(rp, ci, v, X):
    Y = np.zeros((rp.size - 1, X.shape[1]))
    for i in range(rp.size - 1):
        acc = np.zeros(X.shape[1])
        for p in range(rp[i], rp[i + 1]):
            acc = acc + v[p] * X[ci[p]]
        Y[i] = acc
    return Y


def _exchange(dist, rank, recv_list, send_list):
    """grouped point-to-point exchange, as the ncclSend/ncclRecv group of the product"""
    reqs = []
    for p in range(len(send_list)):
        if p == rank:
            continue
        if send_list[p].numel():
            reqs.append(dist.isend(send_list[p].contiguous(), p))
        if recv_list[p].numel():
            reqs.append(dist.irecv(recv_list[p], p))
    for r in reqs:
        r.wait()


def _worker(rank, world, port, kind, g, dim, q):
    try:
        sys.path.insert(0, ROOT)
        import torch
        import torch.distributed as dist
        from rails_b200 import api, problems
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        lib = api.cuda()
        m = g ** dim
        r0, r1 = problems.partition_rows(m, world, rank)
        rows = r1 - r0
        rp, ci, v = problems.build(kind, g, dim, r0, r1)
        nnz = int(rp[-1])
        lcol = np.zeros(max(nnz, 1), dtype=np.int32)
        gg = np.zeros(max(nnz, 1), dtype=np.int32)
        gl = np.zeros(max(nnz, 1), dtype=np.int32)
        need = np.zeros(world, dtype=np.int32)
        ng = ctypes.c_int(0)
        rc = lib.rails_csr_halo_plan(rows, m, r0, world, api.iptr(rp), api.iptr(ci), api.iptr(lcol),
                                     ctypes.byref(ng), api.iptr(gg), api.iptr(gl), api.iptr(need))
        assert rc == 0
        ng = ng.value
        gg, gl = gg[:ng], gl[:ng]
        # plan invariants
        assert need.sum() == ng and need[rank] == 0
        assert np.all(np.diff(gg) > 0)
        owned = (ci >= r0) & (ci < r1)
        assert np.array_equal(lcol[:nnz][owned], ci[owned] - r0)
        assert np.array_equal(gg[lcol[:nnz][~owned] - rows], ci[~owned])
        off = 0
        for p in range(world):
            b, e = problems.partition_rows(m, world, p)
            seg = gg[off:off + need[p]]
            assert np.all((seg >= b) & (seg < e))
            assert np.array_equal(gl[off:off + need[p]], seg - b)
            off += need[p]

        # the exchange rails_csr_create does with ncclAllGather + Send/Recv
        counts = [torch.zeros(world, dtype=torch.int32) for _ in range(world)]
        dist.all_gather(counts, torch.from_numpy(need.copy()))
        send_count = [int(counts[p][rank]) for p in range(world)]
        ids_out = list(torch.from_numpy(gl.copy()).split([int(x) for x in need]))
        ids_in = [torch.zeros(c, dtype=torch.int32) for c in send_count]
        _exchange(dist, rank, ids_in, ids_out)
        send_rows = torch.cat(ids_in).numpy() if world > 1 else np.zeros(0, dtype=np.int32)

        # one halo exchange + local apply, against the global apply
        e_cols = 3
        Xg = np.random.RandomState(7).uniform(-1, 1, (m, e_cols))
        Xl = Xg[r0:r1]
        packed = torch.from_numpy(np.ascontiguousarray(Xl[send_rows]))
        ghost_in = [torch.zeros((int(c), e_cols), dtype=torch.float64) for c in need]
        _exchange(dist, rank, ghost_in, list(packed.split(send_count)))
        ghost = torch.cat(ghost_in).numpy()
        assert np.array_equal(ghost, Xg[gg])
        Y = _csr_apply(rp, lcol, v, np.vstack([Xl, ghost]))
        rpg, cig, vg = problems.build(kind, g, dim)
        Yg = _csr_apply(rpg[r0:r1 + 1] - rpg[r0], cig[rpg[r0]:], vg[rpg[r0]:], Xg)
        assert np.array_equal(Y, Yg)

        # the transposed apply on the partition (a2), as spmm.cu does it: the local
        # block's transpose applied to the owned rows gives rows + nghost values; the
        # ghost part goes back to the owners over the SAME plan with the lists
        # exchanged (what was received is sent, what was sent is received) and is
        # added there peer by peer in rank order.  Against the global A'X.
        Z = np.zeros((rows + ng, e_cols))
        for i in range(rows):
            for pp in range(rp[i], rp[i + 1]):
                Z[lcol[pp]] += v[pp] * Xl[i]
        Yt = Z[:rows].copy()
        back_out = list(torch.from_numpy(np.ascontiguousarray(Z[rows:])).split([int(c) for c in need]))
        back_in = [torch.zeros((c, e_cols), dtype=torch.float64) for c in send_count]
        _exchange(dist, rank, back_in, back_out)
        off = 0
        for p in range(world):
            for j in range(send_count[p]):
                Yt[send_rows[off + j]] += back_in[p][j].numpy()
            off += send_count[p]
        import scipy.sparse as sp
        Ag = sp.csr_matrix((vg, cig, rpg), shape=(m, m))
        want = (Ag.T @ Xg)[r0:r1]
        assert np.max(np.abs(Yt - want)) <= 1e-13 * max(np.max(np.abs(want)), 1.0)
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, "ok"))
    except Exception as exc:           # noqa: BLE001
        import traceback
        q.put((rank, traceback.format_exc() + repr(exc)))


@pytest.mark.parametrize("world,kind,g,dim", [(2, "laplacian", 9, 2), (2, "convdiff", 6, 3),
                                             (3, "laplacian", 5, 3)])
def test_halo_plan_and_exchange(world, kind, g, dim):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, kind, g, dim, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(60)
    for rank, msg in results:
        assert msg == "ok", "rank %d: %s" % (rank, msg)


@pytest.mark.parametrize("case", ["grid", "random", "all_ghost"])
def test_device_plan_algorithm_in_numpy_equals_the_host_plan(case):
    """The steps rails_csr_create runs on the device (spmm.cu: mark ghost columns in a byte
    map, count per 2048-entry chunk, scan the counts, write the flagged indices in order,
    renumber with a binary search in that list), restated in numpy, against the host plan
    rails_csr_halo_plan: same renumbering and ghost list.  (The CUDA kernels themselves are
    compared bit for bit on the GPU, tests/test_gpu_kernels.py.)"""
    sys.path.insert(0, ROOT)
    from rails_b200 import api, problems
    lib = api.cuda()
    rs = np.random.RandomState(3)
    world = 3
    if case == "grid":
        g = 11
        m = g ** 3
        r0, r1 = problems.partition_rows(m, world, 1)
        rp, ci, _ = problems.build("convdiff", g, 3, r0, r1)
    elif case == "random":
        m = 9001
        r0, r1 = problems.partition_rows(m, world, 2)
        per = rs.randint(0, 7, r1 - r0)
        rp = np.concatenate([[0], np.cumsum(per)]).astype(np.int32)
        ci = rs.randint(0, m, int(rp[-1])).astype(np.int32)
    else:
        m = 5000
        r0, r1 = problems.partition_rows(m, world, 1)
        rp = np.arange(r1 - r0 + 1, dtype=np.int32)
        ci = rs.randint(0, r0, r1 - r0).astype(np.int32)
    rows, nnz = r1 - r0, int(rp[-1])
    # host plan
    lcol = np.zeros(max(nnz, 1), dtype=np.int32)
    gg = np.zeros(max(nnz, 1), dtype=np.int32)
    gl = np.zeros(max(nnz, 1), dtype=np.int32)
    need = np.zeros(world, dtype=np.int32)
    ng = ctypes.c_int(0)
    assert lib.rails_csr_halo_plan(rows, m, r0, world, api.iptr(rp), api.iptr(ci), api.iptr(lcol), ctypes.byref(ng),
                                   api.iptr(gg), api.iptr(gl), api.iptr(need)) == 0
    # the device algorithm, step by step
    CH = 2048
    nchunks = -(-m // CH)
    flags = np.zeros(nchunks * CH, dtype=np.uint8)
    ghost_entry = (ci < r0) | (ci >= r1)
    flags[ci[ghost_entry]] = 1                                        # mark_ghost_kernel
    counts = flags.reshape(nchunks, CH).sum(axis=1).astype(np.int64)  # count_ghost_kernel
    offsets = np.concatenate([[0], np.cumsum(counts)])                # host scan of the chunk counts
    glist = np.zeros(int(offsets[-1]), dtype=np.int32)
    for ch in range(nchunks):                                         # write_ghost_kernel: ordered inside a chunk
        idx = np.nonzero(flags[ch * CH:(ch + 1) * CH])[0] + ch * CH
        glist[offsets[ch]:offsets[ch + 1]] = idx
    new = np.where(ghost_entry, rows + np.searchsorted(glist, ci), ci - r0).astype(np.int32)   # renumber_kernel
    assert ng.value == glist.size
    assert np.array_equal(glist, gg[:ng.value])
    assert np.array_equal(new, lcol[:nnz])
