"""GPU parity tests of the SpGEMM path (through the C ABI) against the oracle.

Bar: indptr / indices bit-exact; values BIT-EXACT too (stronger than the 1e-6 relative
tolerance north_star allows), because the kernels accumulate every entry in scipy's order with
unfused multiply and add.  The grid is the reference's own 24-case corpus
(cupy_cusparse/run_all_alg1.sh:5-7); fixtures are test_csr.py:25-135.
"""
import numpy as np
import pytest
import scipy.sparse as sp

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
    c = b200.cusparse.spgemm(a, b, **kw)
    assert c.shape == (A.shape[0], B.shape[1])
    assert c.has_canonical_format
    return c.indptr.cpu().numpy(), c.indices.cpu().numpy(), c.data.cpu().numpy()


@pytest.mark.parametrize("n", helpers.GRID_SIZES)
@pytest.mark.parametrize("d", helpers.GRID_DENSITIES)
def test_grid_bitwise_fp32(b200, n, d):
    A, B = helpers.grid_case(n, d)
    assert_csr_bitwise(dev_spgemm(b200, A, B), oracle_spgemm(A, B), f"grid n={n} d={d}")


@pytest.mark.parametrize("n,d", [(64, 0.1), (256, 0.3), (512, 0.1), (1024, 0.01)])
def test_grid_bitwise_fp64(b200, n, d):
    A, B = helpers.grid_case(n, d, dtype=np.float64)
    assert_csr_bitwise(dev_spgemm(b200, A, B), oracle_spgemm(A, B), f"grid64 n={n} d={d}")


@pytest.mark.parametrize("alg,cf", [(1, 0.2), (2, 0.2), (3, 0.2), (3, 0.05), (3, 1.0)])
def test_alg_modes_identical(b200, alg, cf):
    """ALG1 vs ALG2 vs ALG3 (numerical_error/error.py:25-28 measures their max |diff| for
    cuSPARSE; here every mode must give identical bits)."""
    A, B = helpers.grid_case(512, 0.1)
    assert_csr_bitwise(dev_spgemm(b200, A, B, alg=alg, chunk_fraction=cf), oracle_spgemm(A, B), f"alg={alg} cf={cf}")


@pytest.mark.parametrize("shape", [(2, 3, 4), (4, 3, 2), (300, 50, 1000), (1, 1, 1), (7, 2000, 3)])
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_rectangular_alpha(b200, shape, dtype):
    """test_cusparse.py:372-409 (TestSpgemm): shapes (2,3,4), (4,3,2), density 0.5, alpha=0.5."""
    m, n, k = shape
    A = helpers.random_csr(m, k, 0.5, 1, dtype)
    B = helpers.random_csr(k, n, 0.5, 2, dtype)
    assert_csr_bitwise(dev_spgemm(b200, A, B, alpha=0.5), oracle_spgemm(A, B, alpha=0.5), f"shape={shape}")


def test_fixture_mul_csr(b200):
    """test_csr.py:816-820 test_mul_csr: _make @ _make3 with the explicit zero of _make kept."""
    A = sp.csr_matrix((np.array([0, 1, 2, 3], np.float32), np.array([0, 1, 3, 2], np.int32),
                       np.array([0, 2, 3, 4], np.int32)), shape=(3, 4))
    B = sp.csr_matrix((np.array([1, 2, 3, 4, 5], np.float32), np.array([0, 2, 1, 0, 2], np.int32),
                       np.array([0, 1, 3, 3, 5], np.int32)), shape=(4, 3))
    B.sort_indices()
    p, j, x = dev_spgemm(b200, A, B)
    assert p.tolist() == [0, 3, 5, 5]
    assert j.tolist() == [0, 1, 2, 0, 2]
    assert x.tolist() == [0.0, 3.0, 2.0, 8.0, 10.0]


def test_empty_and_degenerate(b200):
    E = sp.csr_matrix((5, 7), dtype=np.float32)
    F = helpers.random_csr(7, 4, 0.5, 3)
    p, j, x = dev_spgemm(b200, E, F)
    assert p.tolist() == [0] * 6 and j.size == 0 and x.size == 0
    G = helpers.random_csr(6, 5, 0.5, 4)
    p, j, x = dev_spgemm(b200, G, sp.csr_matrix((5, 9), dtype=np.float32))
    assert p.tolist() == [0] * 7 and j.size == 0
    # rows of A that point only at empty rows of B, and empty rows in between
    A = sp.csr_matrix((np.ones(4, np.float32), np.array([0, 2, 2, 3], np.int32), np.array([0, 2, 2, 3, 4], np.int32)),
                      shape=(4, 4))
    B = sp.csr_matrix((np.array([2, 3], np.float32), np.array([1, 5], np.int32), np.array([0, 0, 0, 2, 2], np.int32)),
                      shape=(4, 6))
    assert_csr_bitwise(dev_spgemm(b200, A, B), oracle_spgemm(A, B), "ragged")


def test_cancellation_keeps_explicit_zero(b200):
    A = sp.csr_matrix(np.array([[1.0, 1.0]], np.float32))
    B = sp.csr_matrix(np.array([[2.0, 1.0], [-2.0, 1.0]], np.float32))
    p, j, x = dev_spgemm(b200, A, B)
    assert p.tolist() == [0, 2] and j.tolist() == [0, 1] and x.tolist() == [0.0, 2.0]


@pytest.mark.parametrize("dt", [np.float32, np.float64])
@pytest.mark.parametrize("n,d", [(64, 0.3), (700, 0.2), (2048, 0.4)])
def test_signed_zero_products_and_cancellation(b200, n, d, dt):
    """Explicit zeros (+0.0 and -0.0) in A and B, signed values, exact cancellation: every bin
    (incl. the dense accumulators, which mark untouched columns with a reserved bit pattern) must
    keep the structural entries and give scipy's sign of zero (sums start at +0.0)."""
    rng = np.random.default_rng(n)
    A = helpers.random_csr(n, n, d, 31, dtype=dt)
    B = helpers.random_csr(n, n, d, 32, dtype=dt)
    for M in (A, B):
        M.data[:] = rng.integers(-2, 3, M.nnz).astype(dt)          # small integers: many exact cancellations
        M.data[rng.random(M.nnz) < 0.15] = dt(0.0)
        M.data[rng.random(M.nnz) < 0.15] = dt(-0.0)
    # a row of A made only of zeros against negative B values: all products are -0.0
    A.data[A.indptr[3]:A.indptr[4]] = dt(0.0)
    B.data[rng.random(B.nnz) < 0.3] *= dt(-1)
    got = dev_spgemm(b200, A, B)
    want = oracle_spgemm(A, B)
    assert_csr_bitwise(got, want, f"signed zeros n={n}")
    # scipy drops the entries that are exactly zero (csr_matmat: `if (sums[head] != 0)`);
    # cuSPARSE and the product keep the structural pattern
    C = (A @ B).tocsr()
    C.sort_indices()
    from oracle import oracle
    assert_csr_bitwise(oracle.drop_zero_entries(got), (C.indptr, C.indices, C.data), "filtered vs scipy")
    assert got[1].size > C.nnz


def test_invalid_arguments(b200):
    """test_cusparse.py:413-453: TypeError for non-CSR operands, ValueError for bad shapes."""
    A = b200.csr_matrix(helpers.random_csr(4, 3, 0.5, 1))
    with pytest.raises(TypeError):
        b200.cusparse.spgemm(A, helpers.random_csr(3, 4, 0.5, 2))
    with pytest.raises(ValueError):
        b200.cusparse.spgemm(A, b200.csr_matrix(helpers.random_csr(5, 3, 0.5, 2)))
    with pytest.raises(ValueError):
        b200.cusparse.spgemm(A, b200.csr_matrix(helpers.random_csr(3, 4, 0.5, 2)), alg=3, chunk_fraction=0.0)
    U = b200.csr_matrix((np.array([1, 2], np.float32), np.array([2, 0], np.int32), np.array([0, 2], np.int32)),
                        shape=(1, 3))
    assert not U.has_canonical_format
    with pytest.raises(AssertionError):
        b200.cusparse.spgemm(U, b200.csr_matrix(helpers.random_csr(3, 4, 0.5, 2)))


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_all_bins_large_n(b200, dtype):
    """Row sizes from 1 to ~80 K output entries over n = 300 000 columns: the group, warp,
    CTA-hash and global bitmap-rank kernels all run."""
    A, B = helpers.skewed_pair(dtype=dtype)
    got = dev_spgemm(b200, A, B)
    assert_csr_bitwise(got, oracle_spgemm(A, B), "skewed")
    from spmm_b200 import _lib
    s = _lib.stats(0)
    used = [i for i in range(s.num_bins) if s.num_bin_rows[i] > 0]
    assert len(used) >= 5, f"numeric bins hit: {used}"


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("alg,cf", [(1, 0.2), (3, 0.3)])
def test_heavy_rows_hubs_and_hot_columns(b200, dtype, alg, cf):
    """Column-blocked heavy-row kernel (n = 200 000): hub rows of B give long segments (one piece
    per barrier), the tail rows a few entries per block (ordered windows), and columns 0-3 sit in
    every B row, so each heavy output row sums thousands of products into the same accumulators -
    any slip in the A-row order of those sums changes the low bits."""
    A, B = helpers.hub_pair(dtype=dtype)
    want = oracle_spgemm(A, B)
    got = dev_spgemm(b200, A, B, alg=alg, chunk_fraction=cf)
    assert_csr_bitwise(got, want, "hubs")
    assert_csr_bitwise(dev_spgemm(b200, A, B, alg=alg, chunk_fraction=cf), got, "hubs rerun")


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("mode,acc_kb,tmin", [("rank", None, None), ("rank", "8", None), ("rank", "16", "512"),
                                             ("blocked", None, None), ("mixed", "8", None), ("warp", None, None),
                                             ("warp", None, "512")])
def test_heavy_row_kernels_all_modes(b200, dtype, mode, acc_kb, tmin, monkeypatch):
    """The heavy rows of wide matrices through every kernel choice: the compact rank accumulator
    (k_num_rank_cta; default), forced into many passes per row by a small accumulator (8 KB = 2048
    fp32 ranks per pass against rows of up to 80 K entries: the windowed B-row segments and the
    pass boundaries), the dense column blocks of round 1 (k_num_dense_pieces) and the mix of both.
    All must give scipy's bits."""
    monkeypatch.setenv("B200SP_HEAVY", mode)
    if acc_kb:
        monkeypatch.setenv("B200SP_RANK_ACC_KB", acc_kb)
    if tmin:
        monkeypatch.setenv("B200SP_HEAVY_MIN", tmin)
    if mode == "mixed":
        monkeypatch.setenv("B200SP_HEAVY_DENSE_DIV", "16")
    for name, (A, B) in {"hubs": helpers.hub_pair(dtype=dtype), "skewed": helpers.skewed_pair(dtype=dtype)}.items():
        want = oracle_spgemm(A, B)
        got = dev_spgemm(b200, A, B)
        assert_csr_bitwise(got, want, f"{name} {mode} acc={acc_kb}")
        assert_csr_bitwise(dev_spgemm(b200, A, B, alpha=0.5), oracle_spgemm(A, B, alpha=0.5), f"{name} {mode} alpha")


@pytest.mark.parametrize("turn,flat", [("1", None), ("8", None), ("4", "128"), ("8", "4096")])
def test_blocked_kernel_window_turns(b200, turn, flat, monkeypatch):
    """The short-segment path of k_num_dense_pieces with 1 / 4 / 8 windows per hand-off turn and with
    chunks of long segments forced onto it (B200SP_FLAT_BELOW): same bits whatever the schedule."""
    monkeypatch.setenv("B200SP_HEAVY", "blocked")
    monkeypatch.setenv("B200SP_WINDOW_TURN", turn)
    if flat:
        monkeypatch.setenv("B200SP_FLAT_BELOW", flat)
    for name, (A, B) in {"hubs": helpers.hub_pair(), "skewed": helpers.skewed_pair()}.items():
        assert_csr_bitwise(dev_spgemm(b200, A, B), oracle_spgemm(A, B), f"{name} turn={turn} flat={flat}")
    A, B = helpers.hub_pair(dtype=np.float64, seed=4)
    assert_csr_bitwise(dev_spgemm(b200, A, B), oracle_spgemm(A, B), f"hubs fp64 turn={turn} flat={flat}")


def test_rmat16_bitwise(b200):
    """R-MAT scale 16 A @ A (4.0e8 products, nnz(C) 1.6e8, n = 65 536 > one accumulator): the C5
    recipe at the largest scale the oracle finishes in seconds."""
    A = helpers.rmat_matrix(16)
    want = oracle_spgemm(A, A)
    got = dev_spgemm(b200, A, A)
    assert_csr_bitwise(got, want, "rmat16")


def test_all_bins_chunked(b200):
    A, B = helpers.skewed_pair(seed=11)
    assert_csr_bitwise(dev_spgemm(b200, A, B, alg=3, chunk_fraction=0.3), oracle_spgemm(A, B), "skewed alg3")


def test_config_c1_aliasing_and_determinism(b200):
    """C1: 4096^2, d=1e-3, A @ A with B aliasing A (SpGEMM_vs_SpMV/profiler.py:397-398); twice-run
    equality is the reference's deterministic/test_deterministic.sh."""
    A = helpers.profiler_matrix(4096, 4096, 1e-3)
    want = oracle_spgemm(A, A)
    got1 = dev_spgemm(b200, A, A)
    got2 = dev_spgemm(b200, A, A)
    assert_csr_bitwise(got1, want, "C1")
    assert_csr_bitwise(got2, got1, "C1 rerun")


@pytest.mark.parametrize("d", [1e-4, 1e-3, 1e-2])
def test_config_c3(b200, d):
    """C3: 16384^2 A @ A at the sweep's densities, alg 1 and 3."""
    A = helpers.profiler_matrix(16384, 16384, d)
    want = oracle_spgemm(A, A)
    assert_csr_bitwise(dev_spgemm(b200, A, A), want, f"C3 d={d}")
    assert_csr_bitwise(dev_spgemm(b200, A, A, alg=3), want, f"C3 d={d} alg3")


def test_host_entry_point(b200):
    """b200sp_spgemm_host: host CSR in, host CSR out (the e2e call of bench.py)."""
    A, B = helpers.grid_case(256, 0.1)
    p, j, x, prods = b200.host.spgemm_host(A, B, (256, 256, 256))
    from oracle import oracle
    assert prods == oracle.num_products(A.indptr, A.indices, B.indptr)
    assert_csr_bitwise((p, j, x), oracle_spgemm(A, B), "host")
    p, j, x, _ = b200.host.spgemm_host(A, A, (256, 256, 256), alg=3, chunk_fraction=0.5, alpha=2.0)
    assert_csr_bitwise((p, j, x), oracle_spgemm(A, A, alpha=2.0), "host alias")


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_host_alg3_trades_memory_for_chunks(b200, dtype):
    """ALG3 through the host entry: identical bits, and the device high-water mark (inputs +
    output staging + workspace, on a fresh arena) drops from |C| to two chunk buffers
    (spgemm_from_txt_alg3.cu:101-115 chunk_fraction semantics; SpGEMM_alg_comparison's peak-VRAM axis)."""
    from spmm_b200 import _lib
    A = helpers.profiler_matrix(2048, 2048, 0.02, dtype=dtype)   # 1.7e6 products, nnz(C) ~ 1.1e6
    want = oracle_spgemm(A, A)
    lib, h = _lib.load(), _lib.handle(0)
    peaks, chunks = {}, {}
    for alg, cf in ((1, 0.2), (2, 0.2), (3, 0.5), (3, 0.2), (3, 0.05), (3, 1.0)):
        lib.b200sp_release_workspace(h)
        lib.b200sp_reset_stats(h)
        p, j, x, prods = b200.host.spgemm_host(A, A, A.shape + (A.shape[1],), alg=alg, chunk_fraction=cf)
        assert_csr_bitwise((p, j, x), want, f"host alg={alg} cf={cf}")
        st = _lib.stats(0)
        assert prods == st.num_products == int(np.diff(A.indptr)[A.indices].sum())
        assert st.nnz_c == want[1].size
        peaks[(alg, cf)], chunks[(alg, cf)] = st.peak_workspace_bytes, st.num_chunks
    out_bytes = want[1].size * (4 + np.dtype(dtype).itemsize)
    assert peaks[(2, 0.2)] == peaks[(1, 0.2)] == peaks[(3, 1.0)]
    assert chunks[(3, 0.2)] >= 5 and chunks[(3, 0.05)] >= 20
    # two chunk buffers instead of the whole of C
    assert peaks[(3, 0.2)] <= peaks[(1, 0.2)] - int(0.55 * out_bytes)
    assert peaks[(3, 0.05)] < peaks[(3, 0.2)] < peaks[(3, 0.5)] <= peaks[(1, 0.2)] + 64 * 1024
    # one dominant row: it becomes a chunk of its own, nothing is lost
    S, T = helpers.skewed_pair(k=4500, n=30000)
    p, j, x, _ = b200.host.spgemm_host(S, T, (S.shape[0], S.shape[1], T.shape[1]), alg=3, chunk_fraction=0.01)
    assert_csr_bitwise((p, j, x), oracle_spgemm(S, T), "host alg3 skewed")


def test_noncanonical_b_is_rejected(b200):
    """The C ABI and the host entry refuse a B with unsorted / duplicate columns (the kernels
    rely on distinct columns inside a B row) instead of returning a silently wrong C."""
    A, B = helpers.grid_case(64, 0.3)
    bad = B.copy()
    r = int(np.argmax(np.diff(bad.indptr) >= 2))
    s0 = bad.indptr[r]
    bad.indices[s0], bad.indices[s0 + 1] = bad.indices[s0 + 1], bad.indices[s0]   # one unsorted row
    with pytest.raises(ValueError):
        b200.host.spgemm_host(A, (bad.indptr, bad.indices, bad.data), (64, 64, 64))
    dup = B.copy()
    dup.indices[s0 + 1] = dup.indices[s0]                                         # a duplicate column
    with pytest.raises(ValueError):
        b200.host.spgemm_host(A, (dup.indptr, dup.indices, dup.data), (64, 64, 64))
    p, j, x, _ = b200.host.spgemm_host(A, B, (64, 64, 64))                        # the handle is still usable
    assert_csr_bitwise((p, j, x), oracle_spgemm(A, B), "after rejection")


def test_numeric_reuse_new_values(b200):
    """Same structure, new values: numeric may be called again on the stored plan."""
    import ctypes
    import torch
    from spmm_b200 import _lib
    A, B = helpers.grid_case(128, 0.1)
    a, b = b200.csr_matrix(A), b200.csr_matrix(B)
    plan = b200.cusparse.SpgemmPlan(a, b)  # the two-pass plan (spgemm() may take the single pass, which keeps none)
    c = plan.numeric()
    a2 = b200.csr_matrix((a.data * 3.0, a.indices, a.indptr), shape=a.shape)
    h = plan.h
    cx = torch.empty_like(c.data)
    cj = torch.empty_like(c.indices)
    _lib.check(h, _lib.load().b200sp_spgemm_numeric(h, _lib.R_32F, None, a2.data.data_ptr(), b.data.data_ptr(),
                                                    cj.data_ptr(), cx.data_ptr()))
    A2 = A.copy()
    A2.data = (A.data * np.float32(3.0)).astype(np.float32)
    want = oracle_spgemm(A2, B)
    assert_csr_bitwise((c.indptr.cpu().numpy(), cj.cpu().numpy(), cx.cpu().numpy()), want, "reuse")


def test_streamed_sub_panels(b200):
    """Products with more entries than an int32 indptr can address stream through the GPU in
    row sub-panels (multi.spgemm_panels); here the limit is lowered so that many panels form."""
    from spmm_b200 import multi
    A, B = helpers.skewed_pair()
    b = b200.csr_matrix(B)
    want = helpers.oracle_spgemm(A, B, nthreads=0)
    panels, seen = [], 0
    for p0, p1, c in multi.spgemm_panels(helpers.triple(A), b, (A.shape[0], A.shape[1], B.shape[1]), limit=60000):
        assert p0 == seen and c.shape[0] == p1 - p0 and (c.nnz <= 60000 or p1 - p0 == 1)
        seen = p1
        panels.append((c.indptr.cpu().numpy(), c.indices.cpu().numpy(), c.data.cpu().numpy()))
    assert seen == A.shape[0] and len(panels) >= 5
    helpers.assert_csr_bitwise(multi.concat_panels(panels), want, "streamed panels")
