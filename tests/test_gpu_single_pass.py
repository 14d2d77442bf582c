"""GPU parity tests of the single-pass SpGEMM (b200sp_spgemm_estimate + b200sp_spgemm_single_pass:
no structure pass, C allocated at the bound sum_i min(products_i, n), row offsets by decoupled
look-back) against the oracle and against the two-pass path.  Bar: indptr / indices / values
bit-exact, as for every SpGEMM kernel (tests/test_gpu_spgemm.py).

B200SP_SINGLE_PASS_RATIO=0 makes the library choose the single pass whenever it applies,
B200SP_SINGLE_PASS=0 never.
"""
import ctypes

import numpy as np
import pytest
import scipy.sparse as sp
import torch

import helpers
from helpers import assert_csr_bitwise, oracle_spgemm

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def b200():
    import spmm_b200
    return spmm_b200


def dev_spgemm(b200, A, B, **kw):
    a = b200.csr_matrix(A)
    b = a if B is A else b200.csr_matrix(B)
    a._canonical = b._canonical = True  # by construction; keeps the has_canonical_format kernels out of the launch counts
    c = b200.cusparse.spgemm(a, b, **kw)
    assert c.has_canonical_format
    return c.indptr.cpu().numpy(), c.indices.cpu().numpy(), c.data.cpu().numpy()


def launches_of(b200, fn):
    from spmm_b200 import _lib
    l0 = _lib.stats(0).kernel_launches
    out = fn()
    return out, _lib.stats(0).kernel_launches - l0


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("n,d", [(512, 0.1), (1024, 0.3), (4096, 0.02), (16384, 0.006)])
def test_well_filled_products_take_the_single_pass(b200, n, d, dtype, monkeypatch):
    """products >= m n / 4: the default mode runs row products + canonical check + bins, then ONE kernel."""
    monkeypatch.delenv("B200SP_SINGLE_PASS", raising=False)
    monkeypatch.delenv("B200SP_SINGLE_PASS_RATIO", raising=False)
    m = min(n, 1500)
    A = helpers.profiler_matrix(m, n, d, seed=n, dtype=dtype)
    B = helpers.profiler_matrix(n, n, d, seed=n + 1, dtype=dtype)
    got, nl = launches_of(b200, lambda: dev_spgemm(b200, A, B))
    # the sizes alone say "well filled" here, so the operator takes the direct form: no row lists, no host
    # synchronisation before the kernel (k_row_products, canonical check, k_pack_b, dense)
    assert nl == 4, f"{nl} kernels: not b200sp_spgemm_single_pass_direct"
    want = oracle_spgemm(A, B)
    assert_csr_bitwise(got, want, f"single pass direct n={n} d={d}")
    monkeypatch.setenv("B200SP_SINGLE_PASS_DIRECT", "0")
    got, nl = launches_of(b200, lambda: dev_spgemm(b200, A, B))
    assert nl == 5, f"{nl} kernels: not estimate + single pass (k_row_products, canonical check, scatter, k_pack_b, dense)"
    assert_csr_bitwise(got, want, f"single pass n={n} d={d}")
    monkeypatch.setenv("B200SP_SINGLE_PASS", "0")
    two, nl2 = launches_of(b200, lambda: dev_spgemm(b200, A, B))
    assert nl2 > 5
    assert_csr_bitwise(two, want, "two-pass")


def ragged_pair(dtype, seed=5, m=6000, k=900, n=1500):
    """Rows of A from empty to full, runs of empty rows longer than the 32-row look-back window,
    rows that only point at empty rows of B, explicit and signed zeros."""
    rng = np.random.default_rng(seed)
    lens = rng.choice([0, 0, 1, 2, 7, 40, 200, k], size=m, p=[0.25, 0.1, 0.15, 0.1, 0.1, 0.15, 0.1, 0.05])
    lens[100:260] = 0       # 160 empty rows in a row
    lens[3000:3050] = k
    lens[-1] = 0            # the last row (which reports the total) is empty
    rows = np.repeat(np.arange(m), lens)
    cols = np.concatenate([np.sort(rng.choice(k, L, replace=False)) for L in lens if L] or [np.zeros(0, np.int64)])
    A = sp.csr_matrix((rng.integers(-3, 4, rows.size).astype(dtype), (rows, cols)), shape=(m, k))
    B = sp.random(k, n, density=0.05, format="csr", dtype=dtype, random_state=rng,
                  data_rvs=lambda s: rng.integers(-3, 4, s).astype(dtype))
    B = B.tolil()
    B[5:40] = 0             # empty rows of B
    B = B.tocsr()
    B.data[rng.random(B.nnz) < 0.1] = dtype(-0.0)
    A.sort_indices()
    B.sort_indices()
    return A, B


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("alpha", [1, -0.75])
def test_forced_single_pass_ragged_rows(b200, dtype, alpha, monkeypatch):
    monkeypatch.setenv("B200SP_SINGLE_PASS_RATIO", "0")
    monkeypatch.setenv("B200SP_SINGLE_PASS_DIRECT", "0")
    A, B = ragged_pair(dtype)
    want = oracle_spgemm(A, B, alpha=alpha)
    got, nl = launches_of(b200, lambda: dev_spgemm(b200, A, B, alpha=alpha))
    assert nl == 5
    assert_csr_bitwise(got, want, f"ragged alpha={alpha}")
    # the direct form on the same ragged input (capacity m * n, no estimate): called through the C ABI by the
    # operator whenever the sizes say so; here it is made to by the sizes of this pair
    monkeypatch.delenv("B200SP_SINGLE_PASS_DIRECT")
    got, nl = launches_of(b200, lambda: dev_spgemm(b200, A, B, alpha=alpha))
    assert nl in (4, 5)
    assert_csr_bitwise(got, want, f"ragged direct alpha={alpha}")


def test_forced_single_pass_grid_and_empty(b200, monkeypatch):
    monkeypatch.setenv("B200SP_SINGLE_PASS_RATIO", "0")
    monkeypatch.setenv("B200SP_SINGLE_PASS_DIRECT", "0")
    for n in (32, 256, 1024):
        for d in (0.01, 0.3):
            A, B = helpers.grid_case(n, d)
            assert_csr_bitwise(dev_spgemm(b200, A, B), oracle_spgemm(A, B), f"grid n={n} d={d}")
    E = sp.csr_matrix((5, 7), dtype=np.float32)
    F = helpers.random_csr(7, 4, 0.5, 3)
    p, j, x = dev_spgemm(b200, E, F)
    assert p.tolist() == [0] * 6 and j.size == 0 and x.size == 0
    G = helpers.random_csr(6, 5, 0.5, 4)
    p, j, x = dev_spgemm(b200, G, sp.csr_matrix((5, 9), dtype=np.float32))
    assert p.tolist() == [0] * 7 and j.size == 0


def test_many_rows_deterministic(b200, monkeypatch):
    """40 000 rows over every resident CTA: the look-back chain under load, twice, same bits."""
    monkeypatch.delenv("B200SP_SINGLE_PASS", raising=False)
    monkeypatch.delenv("B200SP_SINGLE_PASS_RATIO", raising=False)
    A = helpers.profiler_matrix(40000, 1024, 0.03, seed=9)
    B = helpers.profiler_matrix(1024, 1024, 0.03, seed=10)
    got, nl = launches_of(b200, lambda: dev_spgemm(b200, A, B))
    assert nl in (4, 5)
    assert_csr_bitwise(got, oracle_spgemm(A, B), "many rows")
    again = dev_spgemm(b200, A, B)
    assert_csr_bitwise(again, got, "second run")


def test_c_abi_estimate_contract(b200):
    """bound >= nnz(C); the tail of the arrays stays untouched; estimate -> symbolic(REUSE) -> numeric
    gives the same C; state and capacity errors."""
    from spmm_b200 import _lib as L
    lib = L.load()
    h = L.new_handle(0)
    try:
        A = helpers.profiler_matrix(700, 2048, 0.02, seed=1)
        B = helpers.profiler_matrix(2048, 2048, 0.02, seed=2)
        want = oracle_spgemm(A, B)
        t = lambda x, dt: torch.from_numpy(np.ascontiguousarray(x)).to("cuda", dt)
        ap, aj, ax = t(A.indptr, torch.int32), t(A.indices, torch.int32), t(A.data, torch.float32)
        bp, bj, bx = t(B.indptr, torch.int32), t(B.indices, torch.int32), t(B.data, torch.float32)
        cp = torch.empty(A.shape[0] + 1, dtype=torch.int32, device="cuda")
        prods, bound, single, nnz = ctypes.c_int64(), ctypes.c_int64(), ctypes.c_int(), ctypes.c_int64()
        P = lambda x: ctypes.c_void_p(x.data_ptr())

        def estimate():
            L.check(h, lib.b200sp_spgemm_estimate(h, L.R_32F, A.shape[0], A.shape[1], B.shape[1], A.nnz, B.nnz, P(ap), P(aj),
                                                  P(bp), P(bj), 0, 0.2, P(cp), ctypes.byref(prods), ctypes.byref(bound),
                                                  ctypes.byref(single)))
        estimate()
        assert single.value == 1 and prods.value == int((A.astype(bool).astype(np.int64) @ np.diff(B.indptr)).sum())
        ub = np.minimum(np.asarray(abs(A).astype(bool).astype(np.int64) @ np.diff(B.indptr)).ravel(), B.shape[1]).sum()
        assert bound.value == int(ub) >= want[1].size
        cap = bound.value
        # too small a capacity is refused before anything is written
        cj = torch.full((cap,), -7, dtype=torch.int32, device="cuda")
        cx = torch.full((cap,), float("nan"), dtype=torch.float32, device="cuda")
        st = lib.b200sp_spgemm_single_pass(h, L.R_32F, None, P(ax), P(bx), P(cj), P(cx), cap - 1, ctypes.byref(nnz))
        assert st == L.ERR_INVALID and b"below the bound" in lib.b200sp_last_error(h)
        L.check(h, lib.b200sp_spgemm_single_pass(h, L.R_32F, None, P(ax), P(bx), P(cj), P(cx), cap, ctypes.byref(nnz)))
        torch.cuda.synchronize()
        k = nnz.value
        assert_csr_bitwise((cp.cpu().numpy(), cj[:k].cpu().numpy(), cx[:k].cpu().numpy()), want, "single pass via C ABI")
        assert bool((cj[k:] == -7).all()) and bool(torch.isnan(cx[k:]).all()), "entries beyond nnz(C) were written"
        # the estimate is consumed: a second single pass needs a new one
        assert lib.b200sp_spgemm_single_pass(h, L.R_32F, None, P(ax), P(bx), P(cj), P(cx), cap, ctypes.byref(nnz)) == L.ERR_STATE
        # estimate -> two-pass continuation
        estimate()
        cp2 = cp
        L.check(h, lib.b200sp_spgemm_symbolic(h, A.shape[0], A.shape[1], B.shape[1], A.nnz, B.nnz, P(ap), P(aj), P(bp), P(bj),
                                              L.ALG_REUSE_ESTIMATE, 0.2, ctypes.byref(prods), ctypes.byref(nnz), P(cp2)))
        assert nnz.value == k
        cj2 = torch.empty(k, dtype=torch.int32, device="cuda")
        cx2 = torch.empty(k, dtype=torch.float32, device="cuda")
        L.check(h, lib.b200sp_spgemm_numeric(h, L.R_32F, None, P(ax), P(bx), P(cj2), P(cx2)))
        torch.cuda.synchronize()
        assert_csr_bitwise((cp2.cpu().numpy(), cj2.cpu().numpy(), cx2.cpu().numpy()), want, "estimate + two-pass")
        # REUSE without a pending estimate
        st = lib.b200sp_spgemm_symbolic(h, A.shape[0], A.shape[1], B.shape[1], A.nnz, B.nnz, P(ap), P(aj), P(bp), P(bj),
                                        L.ALG_REUSE_ESTIMATE, 0.2, ctypes.byref(prods), ctypes.byref(nnz), P(cp2))
        assert st == L.ERR_STATE
        # chunked modes never choose the single pass
        L.check(h, lib.b200sp_spgemm_estimate(h, L.R_32F, A.shape[0], A.shape[1], B.shape[1], A.nnz, B.nnz, P(ap), P(aj), P(bp),
                                              P(bj), 3, 0.2, P(cp), ctypes.byref(prods), ctypes.byref(bound), ctypes.byref(single)))
        assert single.value == 0
    finally:
        lib.b200sp_destroy(h)


def test_direct_entry_contract(b200):
    """b200sp_spgemm_single_pass_direct through the C ABI: result, products, untouched tail, capacity below
    m * n refused, non-canonical B reported after the (skipped) kernel with nothing written."""
    from spmm_b200 import _lib as L
    lib = L.load()
    h = L.new_handle(0)
    try:
        A = helpers.profiler_matrix(300, 1024, 0.05, seed=3, dtype=np.float64)
        B = helpers.profiler_matrix(1024, 900, 0.05, seed=4, dtype=np.float64)
        want = oracle_spgemm(A, B, alpha=-2.0)
        t = lambda x, dt: torch.from_numpy(np.ascontiguousarray(x)).to("cuda", dt)
        ap, aj, ax = t(A.indptr, torch.int32), t(A.indices, torch.int32), t(A.data, torch.float64)
        bp, bj, bx = t(B.indptr, torch.int32), t(B.indices, torch.int32), t(B.data, torch.float64)
        m, k, n = A.shape[0], A.shape[1], B.shape[1]
        cap = m * n
        cp = torch.empty(m + 1, dtype=torch.int32, device="cuda")
        cj = torch.full((cap,), -7, dtype=torch.int32, device="cuda")
        cx = torch.full((cap,), float("nan"), dtype=torch.float64, device="cuda")
        P = lambda x: ctypes.c_void_p(x.data_ptr())
        nnz, prods = ctypes.c_int64(), ctypes.c_int64()
        alpha = ctypes.c_double(-2.0)

        def call(bj_, capacity):
            return lib.b200sp_spgemm_single_pass_direct(h, L.R_64F, m, k, n, A.nnz, B.nnz, P(ap), P(aj), P(ax), P(bp), P(bj_), P(bx),
                                                        ctypes.byref(alpha), P(cp), P(cj), P(cx), capacity, ctypes.byref(nnz),
                                                        ctypes.byref(prods))
        assert call(bj, cap - 1) == L.ERR_INVALID
        bad = bj.clone()
        r = int(np.flatnonzero(np.diff(B.indptr) >= 2)[0])
        s0 = int(B.indptr[r])
        bad[s0], bad[s0 + 1] = bj[s0 + 1].item(), bj[s0].item()  # one unsorted row
        assert call(bad, cap) == L.ERR_INVALID and b"not canonical" in lib.b200sp_last_error(h)
        torch.cuda.synchronize()
        assert bool((cj == -7).all()), "the kernel ran on a non-canonical B"
        L.check(h, call(bj, cap))
        torch.cuda.synchronize()
        kk = nnz.value
        assert prods.value == int((A.astype(bool).astype(np.int64) @ np.diff(B.indptr)).sum())
        assert_csr_bitwise((cp.cpu().numpy(), cj[:kk].cpu().numpy(), cx[:kk].cpu().numpy()), want, "direct via C ABI")
        assert bool((cj[kk:] == -7).all()) and bool(torch.isnan(cx[kk:]).all())
    finally:
        lib.b200sp_destroy(h)

