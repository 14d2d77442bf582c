"""Device format conversions and comparisons through the C ABI,
bit-exact against the oracle (oracle/convert_oracle.c) and against scipy where scipy's bits are
well defined; plus the {csr,csc,coo} x {csr,csc,coo} products of SpGEMM_vs_SpMV/profiler.py."""
import numpy as np
import pytest
import scipy.sparse as sp
import torch

import helpers
from oracle import oracle

pytestmark = pytest.mark.gpu


def _np(t):
    return t.cpu().numpy()


def _csr_np(M):
    return _np(M.indptr), _np(M.indices), _np(M.data)


COO_CASES = [
    # (m, n, entries, generator)
    (64, 64, 900, "pairs"), (5, 2000, 3000, "pairs"), (3000, 7, 4000, "pairs"), (1, 1, 1, "pairs"),
    (1, 1, 700, "heavy"), (1, 5000, 20000, "heavy"), (5000, 1, 20000, "heavy"),
    (200, 150, 50000, "heavy"), (70000, 70000, 300000, "heavy"), (300, 300, 4097, "heavy"),
    (1 << 16, 257, 4096, "pairs"),
]


@pytest.mark.parametrize("dt", [np.float32, np.float64])
@pytest.mark.parametrize("m,n,k,kind", COO_CASES)
def test_coo_to_csr_bitwise(m, n, k, kind, dt):
    import spmm_b200
    gen = helpers.coo_with_pairs if kind == "pairs" else helpers.coo_heavy_duplicates
    row, col, val = gen(m, n, k, 17, dt)
    got = spmm_b200.coo_matrix((val, (row, col)), shape=(m, n)).tocsr()
    assert got.has_canonical_format and got.shape == (m, n)
    helpers.assert_csr_bitwise(_csr_np(got), oracle.coo_to_csr(row, col, val, (m, n)), f"coo {m}x{n} {kind}")
    if kind == "pairs":   # scipy's bits are order-independent here
        M = helpers.scipy_canonical_from_coo(row, col, val, (m, n))
        helpers.assert_csr_bitwise(_csr_np(got), (M.indptr, M.indices, M.data), "vs scipy")


@pytest.mark.parametrize("path", ["radix", "bucket"])
def test_coo_to_csr_both_paths(path, monkeypatch):
    """COO -> CSR has two device paths - row buckets + in-row sort (short rows) and the LSD radix sort over
    (row, column) keys (any row length; also the fallback of the first) - with the same contract."""
    import spmm_b200
    monkeypatch.setenv("B200SP_COO_PATH", path)
    for (m, n, k, kind) in [(64, 64, 900, "pairs"), (200, 150, 50000, "heavy"), (70000, 70000, 300000, "heavy"),
                            (3000, 7, 4000, "pairs"), (1, 5000, 20000, "heavy"), (40, 3000, 30000, "heavy")]:
        gen = helpers.coo_with_pairs if kind == "pairs" else helpers.coo_heavy_duplicates
        row, col, val = gen(m, n, k, 23, np.float32)
        got = spmm_b200.coo_matrix((val, (row, col)), shape=(m, n)).tocsr()
        helpers.assert_csr_bitwise(_csr_np(got), oracle.coo_to_csr(row, col, val, (m, n)), f"coo {path} {m}x{n} {kind}")
    # the transposition uses the same two paths (a segment of more than COO_ROW_CAP entries falls back to radix)
    for (m, n, d) in [(300, 200, 0.1), (5000, 40, 0.3), (3, 9000, 0.4), (20000, 20000, 0.0005)]:
        M = sp.random(m, n, density=d, format="csc", dtype=np.float32, random_state=np.random.RandomState(5))
        got = spmm_b200.csc_matrix(M).tocsr()
        helpers.assert_csr_bitwise(_csr_np(got), oracle.transpose_compressed(M.indptr, M.indices, M.data, m), f"csc->csr {path} {m}x{n}")
    # and csr_matrix.sum_duplicates (rows of up to 300 / 3000 entries: in-row sort and radix fallback)
    rng = np.random.default_rng(4)
    for rows, n, top in [(500, 64, 300), (40, 5000, 3000)]:
        lens = rng.integers(0, top, rows)
        indptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
        indices = rng.integers(0, n, indptr[-1]).astype(np.int32)
        data = rng.standard_normal(indptr[-1]).astype(np.float32)
        g = spmm_b200.csr_matrix((data, indices, indptr), shape=(rows, n))
        g.sum_duplicates()
        helpers.assert_csr_bitwise(_csr_np(g), oracle.csr_sum_duplicates((indptr, indices, data), (rows, n)), f"csr dups {path} {rows}x{n}")


def test_coo_long_duplicate_run_and_determinism():
    import spmm_b200
    rng = np.random.default_rng(5)
    k = 200000
    row = np.full(k, 3, dtype=np.int32)
    col = np.full(k, 9, dtype=np.int32)
    col[::1000] = 2
    val = rng.standard_normal(k).astype(np.float32)
    a = spmm_b200.coo_matrix((val, (row, col)), shape=(8, 16)).tocsr()
    helpers.assert_csr_bitwise(_csr_np(a), oracle.coo_to_csr(row, col, val, (8, 16)), "one hot position")
    b = spmm_b200.coo_matrix((val, (row, col)), shape=(8, 16)).tocsr()
    assert torch.equal(a.data, b.data)


def test_coo_c2_recipe():
    """others/spmm/test.py:6-17 at n = 2^18: randint coordinates, COO -> CSR with duplicates summed."""
    import spmm_b200
    n = 1 << 18
    k = int(4e-5 * n * n)
    rng = np.random.default_rng(0)
    row = rng.integers(0, n, k, dtype=np.int32)
    col = rng.integers(0, n, k, dtype=np.int32)
    val = rng.random(k, dtype=np.float32)
    got = spmm_b200.coo_matrix((val, (row, col)), shape=(n, n)).tocsr()
    helpers.assert_csr_bitwise(_csr_np(got), oracle.coo_to_csr(row, col, val, (n, n)), "c2 recipe")
    M = helpers.scipy_canonical_from_coo(row, col, val, (n, n))
    assert np.array_equal(_np(got.indptr), M.indptr) and np.array_equal(_np(got.indices), M.indices)
    assert np.allclose(_np(got.data), M.data, rtol=1e-6, atol=0)


def test_coo_empty_and_invalid():
    import spmm_b200
    e = spmm_b200.coo_matrix((np.zeros(0, np.float32), (np.zeros(0, np.int32), np.zeros(0, np.int32))), shape=(3, 4)).tocsr()
    assert e.nnz == 0 and _np(e.indptr).tolist() == [0, 0, 0, 0]
    bad = spmm_b200.coo_matrix((np.ones(2, np.float32), (np.array([0, 3]), np.array([0, 0]))), shape=(3, 4))
    with pytest.raises(ValueError):
        bad.tocsr()
    bad = spmm_b200.coo_matrix((np.ones(2, np.float32), (np.array([0, 1]), np.array([0, -1]))), shape=(3, 4))
    with pytest.raises(ValueError):
        bad.tocsr()
    with pytest.raises(ValueError):
        spmm_b200.coo_matrix((np.ones(2, np.float32), (np.array([0]), np.array([0, 1]))), shape=(3, 4))


@pytest.mark.parametrize("dt", [np.float32, np.float64])
@pytest.mark.parametrize("m,n,d", [(50, 70, 0.2), (1, 300, 0.5), (300, 1, 0.5), (40, 40, 0.0), (4096, 4096, 1e-3),
                                   (100000, 300, 0.001), (300, 100000, 0.001)])
def test_csc_csr_roundtrip_bitwise(m, n, d, dt):
    import spmm_b200
    M = sp.random(m, n, density=d, format="csc", dtype=dt, random_state=np.random.RandomState(3))
    R = M.tocsr()
    got = spmm_b200.csc_matrix(M).tocsr()
    helpers.assert_csr_bitwise(_csr_np(got), (R.indptr, R.indices, R.data), "csc->csr vs scipy")
    helpers.assert_csr_bitwise(_csr_np(got), oracle.transpose_compressed(M.indptr, M.indices, M.data, m), "vs oracle")
    back = got.tocsc()
    M.sort_indices()
    helpers.assert_csr_bitwise(_csr_np(back), (M.indptr, M.indices, M.data), "csr->csc")
    T = got.T
    RT = R.T.tocsr()
    RT.sort_indices()
    assert T.shape == (n, m)
    helpers.assert_csr_bitwise(_csr_np(T), (RT.indptr, RT.indices, RT.data), "transpose")


def test_csc_keeps_duplicates_in_input_order():
    import spmm_b200
    # column 0 holds row 2 twice, column 1 holds row 2 once: the CSR row 2 must read
    # (0: 1.0), (0: 2.0), (1: 3.0) - nothing summed, nothing re-ordered
    M = spmm_b200.csc_matrix((np.array([1.0, 2.0, 3.0], np.float32), np.array([2, 2, 2]), np.array([0, 2, 3])), shape=(3, 2))
    R = M.tocsr()
    assert _np(R.indptr).tolist() == [0, 0, 0, 3] and _np(R.indices).tolist() == [0, 0, 1]
    assert _np(R.data).tolist() == [1.0, 2.0, 3.0]
    assert not R.has_canonical_format
    R.sum_duplicates()
    assert _np(R.indices).tolist() == [0, 1] and _np(R.data).tolist() == [3.0, 3.0]


def test_csc_invalid():
    import spmm_b200
    with pytest.raises(ValueError):
        spmm_b200.csc_matrix((np.ones(2, np.float32), np.array([0, 5]), np.array([0, 1, 2])), shape=(3, 2)).tocsr()
    with pytest.raises(ValueError):
        spmm_b200.csc_matrix((np.ones(2, np.float32), np.array([0, 1]), np.array([0, 2, 1])), shape=(3, 2)).tocsr()


@pytest.mark.parametrize("dt", [np.float32, np.float64])
def test_csr_sum_duplicates_bitwise(dt):
    import spmm_b200
    # literal fixtures: test_csr.py:_make_duplicate / _make_unordered
    m = spmm_b200.csr_matrix((np.array([0, 1, 3, 2, 4, 5], dt), np.array([0, 0, 0, 2, 0, 2]), np.array([0, 3, 6, 6])), shape=(3, 4))
    assert not m.has_canonical_format
    m.sum_duplicates()
    assert m.has_canonical_format and m.nnz == 3
    assert _np(m.indptr).tolist() == [0, 1, 3, 3] and _np(m.indices).tolist() == [0, 0, 2] and _np(m.data).tolist() == [4, 4, 7]
    u = spmm_b200.csr_matrix((np.array([1, 2, 3, 4], dt), np.array([1, 0, 1, 2]), np.array([0, 2, 3, 4])), shape=(3, 4))
    u.sum_duplicates()
    assert _np(u.indices).tolist() == [0, 1, 1, 2] and _np(u.data).tolist() == [2, 1, 3, 4]
    # random CSR with unsorted rows and many-fold duplicates
    rng = np.random.default_rng(2)
    rows, n = 500, 64
    lens = rng.integers(0, 300, rows)
    indptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    indices = rng.integers(0, n, indptr[-1]).astype(np.int32)
    data = rng.standard_normal(indptr[-1]).astype(dt)
    g = spmm_b200.csr_matrix((data, indices, indptr), shape=(rows, n))
    g.sum_duplicates()
    helpers.assert_csr_bitwise(_csr_np(g), oracle.csr_sum_duplicates((indptr, indices, data), (rows, n)), "csr dups")


@pytest.mark.parametrize("dt", [np.float32, np.float64])
def test_sort_indices_long_rows_and_duplicates_stable(dt):
    """csr_matrix.sort_indices() (csrsort + gather, cupyx/cusparse.py:832-866): rows longer than the
    shared-memory sort (an R-MAT hub row of a raw product) and rows with duplicate columns go through
    the stable radix sort - equal columns keep their storage order, so a later sum in storage order
    gives the bits of scipy's sum_duplicates on the same input."""
    import spmm_b200
    rng = np.random.default_rng(3)
    n = 90000
    lens = [60000, 0, 5, 17000, 300, 1, 40000, 16385, 2]
    cols = [rng.permutation(n)[:L] for L in lens]               # unsorted, distinct
    cols[4] = rng.integers(0, 40, lens[4])                      # short row, many duplicates
    cols[6] = rng.integers(0, 5000, lens[6])                    # long row with duplicates
    indices = np.concatenate(cols).astype(np.int32)
    indptr = np.concatenate(([0], np.cumsum(lens))).astype(np.int32)
    data = rng.standard_normal(indices.size).astype(dt)
    M = spmm_b200.csr_matrix((data, indices, indptr), shape=(len(lens), n))
    M.sort_indices()
    gp, gj, gx = _csr_np(M)
    assert np.array_equal(gp, indptr)
    for r in range(len(lens)):
        b, e = indptr[r], indptr[r + 1]
        order = np.argsort(indices[b:e], kind="stable")
        assert np.array_equal(gj[b:e], indices[b:e][order]), f"row {r}: columns"
        assert np.array_equal(gx[b:e].view(np.uint8), data[b:e][order].view(np.uint8)), f"row {r}: values not in stable order"
    # an already sorted, duplicate-free matrix is left alone (no fallback, same arrays)
    S = helpers.profiler_matrix(300, 5000, 0.05, seed=8, dtype=dt)
    T = spmm_b200.csr_matrix(S)
    T.sort_indices()
    helpers.assert_csr_bitwise(_csr_np(T), helpers.triple(S), "sorted input")


@pytest.mark.parametrize("dt", [np.float32, np.float64])
def test_toarray_and_max_abs_diff(dt):
    import spmm_b200
    A = helpers.random_csr(300, 170, 0.1, 1, dtype=dt)
    B = helpers.random_csr(300, 170, 0.12, 2, dtype=dt)
    a, b = spmm_b200.csr_matrix(A), spmm_b200.csr_matrix(B)
    dc = a.toarray()
    assert dc.shape == (300, 170) and dc.is_contiguous()
    assert np.array_equal(_np(dc), A.toarray())
    df = a.toarray(order="F")
    assert df.shape == (300, 170) and df.t().is_contiguous() and np.array_equal(_np(df), A.toarray())
    with pytest.raises(ValueError):
        a.toarray(order="K")
    # duplicates accumulate (scipy's toarray does the same)
    d = spmm_b200.csr_matrix((np.array([1, 2, 4, 8], dt), np.array([1, 1, 1, 0]), np.array([0, 3, 4])), shape=(2, 3))
    assert _np(d.toarray()).tolist() == [[0, 7, 0], [8, 0, 0]]
    want = float(np.abs(A.toarray() - B.toarray()).max())     # numerical_error/error.py:27-28
    assert spmm_b200.max_abs_diff(a, b) == want == oracle.csr_max_abs_diff(helpers.triple(A), helpers.triple(B))
    assert spmm_b200.max_abs_diff(a, a) == 0.0
    assert float((a.toarray() - b.toarray()).abs().max()) == want
    Bn = B.copy()
    Bn.data[5] = np.nan
    assert np.isnan(spmm_b200.max_abs_diff(a, spmm_b200.csr_matrix(Bn)))
    with pytest.raises(ValueError):
        spmm_b200.max_abs_diff(a, spmm_b200.csr_matrix(helpers.random_csr(30, 170, 0.1, 1, dtype=dt)))


def test_numerical_error_study_alg1_vs_alg3():
    """numerical_error/error.py:20-29: alg1 vs alg3 max error (ours: exactly 0 by construction)."""
    import spmm_b200
    from spmm_b200 import cusparse
    for dim, dens in [(256, 0.1), (512, 0.5), (1024, 0.01)]:
        A = spmm_b200.csr_matrix(helpers.random_csr(dim, dim, dens, 10))
        B = spmm_b200.csr_matrix(helpers.random_csr(dim, dim, dens, 11))
        C = cusparse.spgemm(A, B, alg=1)
        D = cusparse.spgemm(A, B, alg=3, chunk_fraction=0.3)
        assert spmm_b200.max_abs_diff(C, D) == 0.0
        assert float((C.toarray() - D.toarray()).abs().max()) == 0.0


@pytest.mark.parametrize("fa", ["csr", "csc", "coo"])
@pytest.mark.parametrize("fb", ["csr", "csc", "coo"])
def test_format_grid_products(fa, fb):
    """SpGEMM_vs_SpMV/profiler.py:404-411: every operand format pair, and A_fmt @ dense vector."""
    import spmm_b200
    mk = {"csr": spmm_b200.csr_matrix, "csc": spmm_b200.csc_matrix, "coo": spmm_b200.coo_matrix}
    A = helpers.profiler_matrix(1024, 1024, 1e-2, seed=42)
    B = helpers.profiler_matrix(1024, 1024, 1e-2, seed=43)
    C = mk[fa](A.asformat(fa)) @ mk[fb](B.asformat(fb))
    assert C.format == "csr"
    helpers.assert_csr_bitwise(_csr_np(C), helpers.oracle_spgemm(A, B), f"{fa}@{fb}")
    if fb == "csr":
        x = np.random.default_rng(44).standard_normal((1024, 1)).astype(np.float32)
        y = mk[fa](A.asformat(fa)) @ torch.from_numpy(x).cuda()
        ref = A @ x
        assert tuple(y.shape) == (1024, 1)
        bound = abs(A) @ np.abs(x)
        assert np.all(np.abs(_np(y) - ref) <= 1e-5 * bound + 1e-30)
