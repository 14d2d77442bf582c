"""CPU tests of the row-panel multi-GPU host logic (spmm_b200/multi.py): partitioning, the
packed B broadcast and the panel concatenation, on the gloo backend with world_size 2.
The per-panel product is injected (the oracle stands in for the GPU kernels here: this test
checks the plumbing, the -m gpu tests check the kernels)."""
import os
import socket

import numpy as np
import pytest
import scipy.sparse as sp
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import helpers
from spmm_b200 import multi


def test_row_products_and_balanced_panels():
    A, B = helpers.skewed_pair(k=5000, n=20000)
    rp = multi.row_products(A.indptr, A.indices, B.indptr)
    from oracle import oracle
    assert np.array_equal(rp, oracle.row_products(A.indptr, A.indices, B.indptr))
    for parts in (1, 2, 3, 8):
        b = multi.balanced_row_panels(rp, parts)
        assert b[0] == 0 and b[-1] == A.shape[0] and len(b) == parts + 1
        assert all(b[i] <= b[i + 1] for i in range(parts))
        w = [int(rp[b[i]:b[i + 1]].sum()) for i in range(parts)]
        # no panel carries more than its fair share plus one (heaviest) row
        assert max(w) <= rp.sum() / parts + rp.max() + A.shape[0]
    # more parts than rows: empty panels are legal
    b = multi.balanced_row_panels(np.array([5, 5]), 4)
    assert b[0] == 0 and b[-1] == 2 and len(b) == 5


def test_block_cyclic_panels_cover_rows_once():
    rng = np.random.default_rng(5)
    w = (rng.pareto(1.2, size=1000) * 50).astype(np.int64)
    for parts, per in ((1, 1), (2, 4), (8, 4), (3, 5)):
        dealt = multi.block_cyclic_panels(w, parts, per)
        assert len(dealt) == parts
        ranges = sorted(r for part in dealt for r in part)
        assert ranges[0][0] == 0 and ranges[-1][1] == len(w)
        assert all(ranges[i][1] == ranges[i + 1][0] for i in range(len(ranges) - 1))  # disjoint, no gap
        assert all(len(part) <= per for part in dealt)
        loads = [sum(int(w[a:b].sum()) for a, b in part) for part in dealt]
        assert max(loads) <= w.sum() / parts + per * (w.max() + 1) + len(w)
    # fewer rows than panels: empty panels are dropped
    assert sum(len(p) for p in multi.block_cyclic_panels(np.array([3, 3]), 4, 4)) <= 2


def test_slice_and_concat_roundtrip():
    M = helpers.random_csr(40, 30, 0.2, 4)
    bounds = [0, 7, 7, 25, 40]
    panels = [multi.slice_rows(M.indptr, M.indices, M.data, bounds[i], bounds[i + 1]) for i in range(4)]
    assert panels[1][0].tolist() == [0]
    p, j, x = multi.concat_panels(panels)
    assert np.array_equal(p, M.indptr) and np.array_equal(j, M.indices) and np.array_equal(x, M.data)
    assert p.dtype == np.int32


def test_pack_unpack_csr():
    M = helpers.random_csr(12, 9, 0.4, 5, dtype=np.float64)
    t = tuple(torch.from_numpy(a) for a in (M.indptr, M.indices, M.data))
    buf = multi.pack_csr(*t)
    assert buf.numel() == 13 * 4 + M.nnz * 12
    p, j, x = multi.unpack_csr(buf, 12, M.nnz, torch.float64)
    assert torch.equal(p, t[0]) and torch.equal(j, t[1]) and torch.equal(x, t[2])


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _oracle_compute(panel, Bd, shape):
    from oracle import oracle
    cp, cj, cx = oracle.spgemm(tuple(t.numpy() for t in panel), tuple(t.numpy() for t in Bd), shape[2])
    return torch.from_numpy(cp), torch.from_numpy(cj), torch.from_numpy(cx)


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        A, B = helpers.skewed_pair(k=4500, n=9000)
        Bt = helpers.triple(B) if rank == 0 else None          # only rank 0 holds B
        out = multi.spgemm_row_panels(helpers.triple(A), Bt, (A.shape[0], A.shape[1], B.shape[1]),
                                      compute=_oracle_compute)
        r0, r1, panel = multi.spgemm_row_panels(helpers.triple(A), Bt, (A.shape[0], A.shape[1], B.shape[1]),
                                                compute=_oracle_compute, gather=False)
        q.put((rank, out, (r0, r1, panel[0][-1])))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_row_panels_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = {}
    for _ in range(2):
        rank, out, info = q.get(timeout=90)
        res[rank] = (out, info)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[1][0] is None
    A, B = helpers.skewed_pair(k=4500, n=9000)
    C = (A @ B).tocsr()
    C.sort_indices()
    helpers.assert_csr_bitwise(res[0][0], helpers.triple(C), "world2 concat")
    (a0, a1, n0), (b0, b1, n1) = res[0][1], res[1][1]
    assert a0 == 0 and a1 == b0 and b1 == A.shape[0] and 0 < a1 < A.shape[0]
    assert int(n0) + int(n1) == C.nnz


def test_split_rows_by_bound():
    rp = np.array([5, 100, 3, 0, 50, 7, 7, 7])
    assert multi.split_rows_by_bound(rp, 60, 0, 8, limit=70) == [0, 4, 7, 8]
    assert multi.split_rows_by_bound(rp, 60, 2, 8, limit=1000) == [2, 8]
    # a limit below one row's bound still makes progress, one row per panel
    assert multi.split_rows_by_bound(rp, 60, 0, 8, limit=10) == [0, 1, 2, 4, 5, 6, 7, 8]
    cuts = multi.split_rows_by_bound(rp, 60, 0, 8, limit=70)
    ub = np.minimum(rp, 60)
    assert all(ub[a:b].sum() <= 70 for a, b in zip(cuts[:-1], cuts[1:]))
    assert multi.split_rows_by_bound(rp, 60, 3, 3) == [3]


def _oracle_spmv(panel, xd, shape):
    from oracle import oracle
    return torch.from_numpy(oracle.spmv(tuple(t.numpy() for t in panel), xd.numpy()))


def _spmv_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        A = helpers.random_csr(301, 170, 0.1, 9)
        A.data[:] = np.arange(A.nnz) % 7 - 3.0          # a few heavy rows shift the split
        x = np.random.default_rng(3).standard_normal(170).astype(np.float32) if rank == 0 else None
        y = multi.spmv_row_panels(helpers.triple(A), x, A.shape, compute=_oracle_spmv)
        r0, r1, yp = multi.spmv_row_panels(helpers.triple(A), x, A.shape, compute=_oracle_spmv, gather=False)
        q.put((rank, y, (r0, r1, yp.size)))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_spmv_row_panels_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_spmv_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = {}
    for _ in range(2):
        rank, y, info = q.get(timeout=90)
        res[rank] = (y, info)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[1][0] is None
    A = helpers.random_csr(301, 170, 0.1, 9)
    A.data[:] = np.arange(A.nnz) % 7 - 3.0
    x = np.random.default_rng(3).standard_normal(170).astype(np.float32)
    from oracle import oracle
    want = oracle.spmv(helpers.triple(A), x)
    assert np.array_equal(res[0][0].view(np.uint32), want.view(np.uint32))   # only rank 0 held x
    (a0, a1, n0), (b0, b1, n1) = res[0][1], res[1][1]
    assert a0 == 0 and a1 == b0 and b1 == 301 and n0 == a1 and n1 == b1 - b0 and 0 < a1 < 301
