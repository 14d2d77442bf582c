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
    # a dominant first row (an R-MAT hub) ends its own panel; the other ranks still get rows
    b = multi.balanced_row_panels(np.array([100, 0, 0, 0, 0, 0, 0, 1]), 4)
    assert b[0] == 0 and b[1] == 1 and b[-1] == 8
    b = multi.balanced_row_panels(np.array([1000, 10, 10, 10, 10, 10, 10, 10, 10]), 2)
    assert b == [0, 1, 9]


def test_host_row_products_native_contract():
    """b200sp_host_row_products (host code of libb200sp, no device): int64 sums, empty rows, any index dtype
    on the Python side, and a clear error for a column index outside B's rows or a malformed indptr."""
    from oracle import oracle
    rng = np.random.default_rng(12)
    A = sp.random(3000, 800, density=0.02, format="csr", dtype=np.float32, random_state=rng)
    B = sp.random(800, 500, density=0.3, format="csr", dtype=np.float32, random_state=rng)
    want = oracle.row_products(A.indptr, A.indices, B.indptr)
    got = multi.row_products(A.indptr.astype(np.int64), A.indices.astype(np.int64), B.indptr.astype(np.int64))
    assert got.dtype == np.int64 and np.array_equal(got, want)
    assert np.array_equal(got, np.asarray(A.astype(bool).astype(np.int64) @ np.diff(B.indptr)).ravel())
    assert multi.row_products(np.zeros(1, np.int32), np.zeros(0, np.int32), B.indptr).size == 0
    bad = A.indices.copy()
    bad[5] = 800
    with pytest.raises(ValueError):
        multi.row_products(A.indptr, bad, B.indptr)
    with pytest.raises(ValueError):
        multi.row_products(np.array([0, 3, 2], np.int32), np.array([0, 1, 2], np.int32), B.indptr)
    # a sum beyond int32: 70 000 references to a row of 40 000 entries
    Bp = np.array([0, 40000], np.int32)
    Ap = np.array([0, 70000], np.int32)
    assert multi.row_products(Ap, np.zeros(70000, np.int32), Bp).tolist() == [70000 * 40000]


def test_cast_csr_host_int64_indices():
    """SciPy hands out int64 index arrays for COO-built matrices: the packed broadcast is sized
    for int32, so the host cast must normalise them (and refuse what does not fit)."""
    M = helpers.random_csr(12, 9, 0.4, 5, dtype=np.float64)
    p, j, x = multi.cast_csr_host((M.indptr.astype(np.int64), M.indices.astype(np.int64), M.data.astype(np.float64)))
    assert p.dtype == np.int32 and j.dtype == np.int32 and x.dtype == np.float64
    assert np.array_equal(p, M.indptr) and np.array_equal(j, M.indices)
    assert multi.cast_csr_host(M, np.float32)[2].dtype == np.float32
    with pytest.raises(OverflowError):
        multi.cast_csr_host((np.array([0, 1], np.int64), np.array([1 << 31], np.int64), np.ones(1, np.float32)))
    with pytest.raises(TypeError):
        multi.pack_csr(torch.zeros(2, dtype=torch.int64), torch.zeros(1, dtype=torch.int32), torch.zeros(1))


def test_plan_sub_panels_cover_rows_once():
    rng = np.random.default_rng(6)
    rp = (rng.pareto(1.1, size=500) * 40).astype(np.int64)
    for world, per, limit in ((1, 1, 10 ** 9), (2, 4, 10 ** 9), (4, 4, 300), (3, 2, 50)):
        subs = multi.plan_sub_panels(rp, 64, world, per, limit)
        assert len(subs) == world
        flat = sorted(s for part in subs for s in part)
        assert flat[0][0] == 0 and flat[-1][1] == len(rp)
        assert all(flat[i][1] == flat[i + 1][0] for i in range(len(flat) - 1))
        ub = np.minimum(rp, 64)
        assert all(ub[a:b].sum() <= limit or b - a == 1 for a, b in flat)


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
        # only rank 0 holds B, and with int64 index arrays (what scipy builds from COO input)
        Bt = (B.indptr.astype(np.int64), B.indices.astype(np.int64), B.data) if rank == 0 else None
        shape = (A.shape[0], A.shape[1], B.shape[1])
        st = {}
        out = multi.spgemm_row_panels(helpers.triple(A), Bt, shape, compute=_oracle_compute, copy=True, stats=st)
        # a small limit forces extra sub-panel cuts; the shared buffer is reused (no growth)
        subs = multi.spgemm_row_panels(helpers.triple(A), Bt, shape, compute=_oracle_compute, gather=False,
                                       panels_per_rank=3, limit=40000)
        info = [(p0, p1, int(c[0][-1]), int(c[1].size), c[0].dtype == np.int32 and int(c[0][0]) == 0) for p0, p1, c in subs]
        q.put((rank, out, info, st))
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
        rank, out, info, st = q.get(timeout=90)
        res[rank] = (out, info, st)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[1][0] is None
    A, B = helpers.skewed_pair(k=4500, n=9000)
    C = (A @ B).tocsr()
    C.sort_indices()
    helpers.assert_csr_bitwise(res[0][0], helpers.triple(C), "world2 concat in the shared host buffer")
    assert res[0][2]["nnz"] == C.nnz and sum(res[0][2]["sub_panels"]) >= 2
    # gather=False: every rank's sub-panels, interleaved (block-cyclic) and covering the rows once
    flat = sorted(res[0][1] + res[1][1])
    assert flat[0][0] == 0 and flat[-1][1] == A.shape[0]
    assert all(flat[i][1] == flat[i + 1][0] for i in range(len(flat) - 1))
    assert all(ok and last == size for _p0, _p1, last, size, ok in flat)
    assert sum(f[3] for f in flat) == C.nnz
    assert len(res[0][1]) >= 2 and len(res[1][1]) >= 2
    ub = np.minimum(multi.row_products(A.indptr, A.indices, B.indptr), B.shape[1])
    assert all(ub[p0:p1].sum() <= 40000 or p1 - p0 == 1 for p0, p1, *_ in flat)


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
    assert np.array_equal(res[1][0], res[0][0])  # y is all-gathered: every rank holds it
    A = helpers.random_csr(301, 170, 0.1, 9)
    A.data[:] = np.arange(A.nnz) % 7 - 3.0
    x = np.random.default_rng(3).standard_normal(170).astype(np.float32)
    from oracle import oracle
    want = oracle.spmv(helpers.triple(A), x)
    assert np.array_equal(res[0][0].view(np.uint32), want.view(np.uint32))   # only rank 0 held x
    (a0, a1, n0), (b0, b1, n1) = res[0][1], res[1][1]
    assert a0 == 0 and a1 == b0 and b1 == 301 and n0 == a1 and n1 == b1 - b0 and 0 < a1 < 301


def _queue_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pq = multi.PanelQueue()
        got = []
        for _round in range(2):  # two jobs over the same queue: reset() brings the counter back to 0
            pq.reset()
            mine = []
            while True:
                i = pq.next()
                if i >= 37:
                    break
                mine.append(i)
            got.append(mine)
        pq.close()
        q.put((rank, got))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_panel_queue_world2_gloo():
    """The dynamic panel queue of the C5 job: every index is handed out exactly once per round, to whichever
    rank asks first; work_ordered_panels covers the rows once, heaviest panel first, within the bound."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_queue_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=90) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rnd in range(2):
        assert sorted(res[0][rnd] + res[1][rnd]) == list(range(37))
    rng = np.random.default_rng(8)
    rp = (rng.pareto(1.1, size=700) * 40).astype(np.int64)
    panels = multi.work_ordered_panels(rp, 64, 12, limit=400)
    flat = sorted(panels)
    assert flat[0][0] == 0 and flat[-1][1] == 700 and all(flat[i][1] == flat[i + 1][0] for i in range(len(flat) - 1))
    w = [int(rp[a:b].sum()) for a, b in panels]
    assert w == sorted(w, reverse=True)
    assert all(np.minimum(rp[a:b], 64).sum() <= 400 or b - a == 1 for a, b in panels)
