"""Pins oracle/convert_oracle.c (the CPU restatement of the format conversions, SURVEY.md
section 8f N2/N4) against the reference's literal fixtures, the committed golden vectors
(scipy, tests/golden/make_golden.py) and live scipy.  CPU only."""
import os

import numpy as np
import pytest
import scipy.sparse as sp

import helpers
from oracle import oracle

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "convert_small.npz")


def _bits(a):
    return a.view(np.uint32 if a.dtype == np.float32 else np.uint64)


@pytest.mark.parametrize("dt", [np.float32, np.float64])
def test_literal_fixtures_of_reference_tests(dt):
    # test_coo.py:_make_duplicate: rows [1,1,1,1,0,1], cols [0,0,2,0,0,2], data 0..5 ->
    # [[4,0,0,0],[4,0,7,0],[0,0,0,0]], nnz 3 (test_sum_duplicates asserts nnz == 3)
    row, col = np.array([1, 1, 1, 1, 0, 1]), np.array([0, 0, 2, 0, 0, 2])
    p, j, x = oracle.coo_to_csr(row, col, np.arange(6, dtype=dt), (3, 4))
    assert p.tolist() == [0, 1, 3, 3] and j.tolist() == [0, 0, 2] and x.tolist() == [4, 4, 7]
    # test_coo.py:_make_unordered -> [[1,2,0,0],[0,0,0,3],[0,0,4,0]]
    p, j, x = oracle.coo_to_csr([0, 2, 1, 0], [0, 2, 3, 1], np.array([1, 4, 3, 2], dtype=dt), (3, 4))
    assert p.tolist() == [0, 2, 3, 4] and j.tolist() == [0, 1, 3, 2] and x.tolist() == [1, 2, 3, 4]
    # test_csr.py:_make_duplicate (CSR with duplicates) -> same matrix, nnz 3
    p, j, x = oracle.csr_sum_duplicates(([0, 3, 6, 6], [0, 0, 0, 2, 0, 2], np.array([0, 1, 3, 2, 4, 5], dtype=dt)), (3, 4))
    assert p.tolist() == [0, 1, 3, 3] and j.tolist() == [0, 0, 2] and x.tolist() == [4, 4, 7]
    # test_csr.py:_make_unordered -> rows sorted: [[2,1,0,0],[0,3,0,0],[0,0,4,0]]
    p, j, x = oracle.csr_sum_duplicates(([0, 2, 3, 4], [1, 0, 1, 2], np.array([1, 2, 3, 4], dtype=dt)), (3, 4))
    assert p.tolist() == [0, 2, 3, 4] and j.tolist() == [0, 1, 1, 2] and x.tolist() == [2, 1, 3, 4]
    # empty (test_coo.py:_make_empty)
    p, j, x = oracle.coo_to_csr([], [], np.array([], dtype=dt), (3, 4))
    assert p.tolist() == [0, 0, 0, 0] and j.size == 0 and x.size == 0


def test_golden_vectors():
    g = np.load(GOLD)
    names = sorted({k.split("/")[0] for k in g.files})
    assert len(names) == 5
    for name in names:
        m, n = (int(v) for v in g[f"{name}/shape"])
        if name.startswith("coo"):
            p, j, x = oracle.coo_to_csr(g[f"{name}/row"], g[f"{name}/col"], g[f"{name}/val"], (m, n))
        else:
            p, j, x = oracle.transpose_compressed(g[f"{name}/ptr"], g[f"{name}/idx"], g[f"{name}/val"], m)
        helpers.assert_csr_bitwise((p, j, x), (g[f"{name}/Cp"], g[f"{name}/Cj"], g[f"{name}/Cx"]), name)


@pytest.mark.parametrize("dt", [np.float32, np.float64])
def test_coo_to_csr_live_scipy(dt):
    # at most two-fold duplicates: bit-identical to scipy
    for (m, n, k, seed) in [(64, 64, 900, 1), (5, 2000, 3000, 2), (3000, 7, 4000, 3), (1, 1, 1, 4)]:
        row, col, val = helpers.coo_with_pairs(m, n, k, seed, dt)
        M = helpers.scipy_canonical_from_coo(row, col, val, (m, n))
        helpers.assert_csr_bitwise(oracle.coo_to_csr(row, col, val, (m, n)),
                                   (M.indptr, M.indices, M.data), f"pairs {m}x{n}")
    # many-fold duplicates: structure identical, values equal up to summation order
    row, col, val = helpers.coo_heavy_duplicates(200, 150, 50000, 9, dt)
    M = helpers.scipy_canonical_from_coo(row, col, val, (200, 150))
    p, j, x = oracle.coo_to_csr(row, col, val, (200, 150))
    assert np.array_equal(p, M.indptr) and np.array_equal(j, M.indices)
    absM = helpers.scipy_canonical_from_coo(row, col, np.abs(val), (200, 150))
    eps = np.finfo(dt).eps
    assert np.all(np.abs(x - M.data) <= 4 * eps * 3000 * absM.data)
    # and the stated rule itself: input-order sequential sums
    lin = row.astype(np.int64) * 150 + col
    hot = (200 // 2) * 150 + 150 // 3
    acc = dt(0)
    first = True
    for v in val[lin == hot]:
        acc = v if first else dt(acc + v)
        first = False
    pos = p[200 // 2] + int(np.searchsorted(j[p[100]:p[101]], 150 // 3))
    assert _bits(np.array([x[pos]]))[0] == _bits(np.array([acc]))[0]


def test_coo_out_of_range():
    with pytest.raises(ValueError):
        oracle.coo_to_csr([0, 3], [0, 0], np.ones(2, np.float32), (3, 4))
    with pytest.raises(ValueError):
        oracle.coo_to_csr([0, 1], [0, -1], np.ones(2, np.float32), (3, 4))


@pytest.mark.parametrize("dt", [np.float32, np.float64])
def test_transpose_live_scipy(dt):
    for (m, n, d, seed) in [(50, 70, 0.2, 1), (1, 300, 0.5, 2), (300, 1, 0.5, 3), (40, 40, 0.0, 4)]:
        M = sp.random(m, n, density=d, format="csc", dtype=dt, random_state=np.random.RandomState(seed))
        R = M.tocsr()
        helpers.assert_csr_bitwise(oracle.transpose_compressed(M.indptr, M.indices, M.data, m),
                                   (R.indptr, R.indices, R.data), "csc->csr")
        back = oracle.transpose_compressed(R.indptr, R.indices, R.data, n)   # csr -> csc
        M.sort_indices()
        helpers.assert_csr_bitwise(back, (M.indptr, M.indices, M.data), "csr->csc")


def test_max_abs_diff_matches_dense_statement():
    # numerical_error/error.py:27-28: abs(C.toarray() - D.toarray()).max()
    for dt in (np.float32, np.float64):
        A = helpers.random_csr(60, 45, 0.2, 1, dtype=dt)
        B = helpers.random_csr(60, 45, 0.25, 2, dtype=dt)
        want = float(np.abs(A.toarray() - B.toarray()).max())
        assert oracle.csr_max_abs_diff(helpers.triple(A), helpers.triple(B)) == want
        assert oracle.csr_max_abs_diff(helpers.triple(A), helpers.triple(A)) == 0.0
    A = helpers.random_csr(10, 10, 0.5, 3)
    B = A.copy()
    B.data[3] = np.nan
    assert np.isnan(oracle.csr_max_abs_diff(helpers.triple(A), helpers.triple(B)))
