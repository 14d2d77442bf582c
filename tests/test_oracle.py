"""Pins the CPU oracle (oracle/spgemm_oracle.c) - CPU only, no GPU needed.

(a) literal known-answer fixtures of the reference's vendored test-suite
    (modify_src/cupy-src/tests/cupyx_tests/scipy_tests/sparse_tests/test_csr.py:25-135, products
    computed by hand below);
(b) the committed golden vectors tests/golden/*.npz + checksums.json, produced by the
    reference's CPU path (scipy.sparse) with tests/golden/make_golden.py - bit for bit;
(c) live scipy on the 24-case grid of cupy_cusparse/run_all_alg1.sh:5-7.
"""
import hashlib
import json
import os

import numpy as np
import pytest
import scipy.sparse as sp

import helpers
from oracle import oracle

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


# ---- (a) literal fixtures ---------------------------------------------------------------

def _make(dt):  # test_csr.py:25-33
    return (np.array([0, 2, 3, 4], np.int32), np.array([0, 1, 3, 2], np.int32), np.array([0, 1, 2, 3], dt)), (3, 4)


def _make3(dt):  # test_csr.py:60-68
    return (np.array([0, 1, 3, 3, 5], np.int32), np.array([0, 2, 1, 0, 2], np.int32), np.array([1, 2, 3, 4, 5], dt)), (4, 3)


def _make4(dt):  # test_csr.py:71-78
    return (np.array([0, 3, 6, 9], np.int32), np.array([0, 2, 3, 0, 1, 3, 0, 1, 2], np.int32),
            np.arange(1, 10).astype(dt)), (3, 4)


def _make_empty(dt):  # test_csr.py:103-107
    return (np.zeros(4, np.int32), np.zeros(0, np.int32), np.zeros(0, dt)), (3, 4)


@pytest.mark.parametrize("dt", [np.float32, np.float64])
def test_fixture_mul_csr_hand_computed(dt):
    """test_csr.py:816-820 test_mul_csr: _make @ _make3.  _make3's row 1 is stored unsorted
    (cols 2, 1) - the product must not care.  The explicit zero A[0,0] gives a stored 0."""
    (A, _), (B, bs) = _make(dt), _make3(dt)
    p, j, x = oracle.spgemm(A, B, bs[1])
    assert p.tolist() == [0, 3, 5, 5]
    assert j.tolist() == [0, 1, 2, 0, 2]
    assert x.tolist() == [0, 3, 2, 8, 10]
    (A4, _) = _make4(dt)
    p, j, x = oracle.spgemm(A4, B, bs[1])
    assert p.tolist() == [0, 2, 5, 8]
    assert j.tolist() == [0, 2, 0, 1, 2, 0, 1, 2]
    assert x.tolist() == [13, 15, 28, 15, 40, 7, 24, 16]
    (E, _) = _make_empty(dt)
    p, j, x = oracle.spgemm(E, B, bs[1])
    assert p.tolist() == [0, 0, 0, 0] and j.size == 0 and x.size == 0


@pytest.mark.parametrize("dt", [np.float32, np.float64])
def test_fixture_mul_dense_vector(dt):
    """test_csr.py:851-855 test_mul_dense_vector: _make @ arange(4) = [1, 6, 6]."""
    (A, _) = _make(dt)
    assert oracle.spmv(A, np.arange(4).astype(dt)).tolist() == [1, 6, 6]
    (A4, _) = _make4(dt)
    assert oracle.spmv(A4, np.arange(4).astype(dt)).tolist() == [13, 23, 26]
    assert oracle.spmv(A4, np.arange(4).astype(dt), y=np.ones(3, dt), alpha=2, beta=3).tolist() == [29, 49, 55]


def test_num_products_and_canonical():
    (A, _), (B, _) = _make(np.float32), _make3(np.float32)
    # rows of B have 1, 2, 0, 2 entries; A's columns are 0,1 | 3 | 2
    assert oracle.num_products(A[0], A[1], B[0]) == 1 + 2 + 2 + 0
    assert oracle.row_products(A[0], A[1], B[0]).tolist() == [3, 2, 0]
    assert oracle.is_canonical(A[0], A[1], 4)
    assert not oracle.is_canonical(B[0], B[1], 3)        # row 1 unsorted (test_csr.py:63)
    dup = (np.array([0, 3, 6, 6], np.int32), np.array([0, 0, 0, 2, 0, 2], np.int32))  # _make_duplicate :92-100
    assert not oracle.is_canonical(dup[0], dup[1], 4)
    assert not oracle.is_canonical(A[0], A[1], 3)        # column 3 out of range


# ---- (b) committed golden vectors -------------------------------------------------------

def _cases(npz):
    return sorted({k.split("/")[0] for k in npz.files})


def test_golden_spgemm_bitwise():
    g = np.load(os.path.join(GOLD, "spgemm_small.npz"))
    names = _cases(g)
    assert len(names) >= 8
    for name in names:
        A = (g[f"{name}/Ap"], g[f"{name}/Aj"], g[f"{name}/Ax"])
        B = (g[f"{name}/Bp"], g[f"{name}/Bj"], g[f"{name}/Bx"])
        n = int(g[f"{name}/shape"][2])
        got = oracle.spgemm(A, B, n, alpha=float(g[f"{name}/alpha"]))
        helpers.assert_csr_bitwise(got, (g[f"{name}/Cp"], g[f"{name}/Cj"], g[f"{name}/Cx"]), name)


def test_golden_spmv():
    g = np.load(os.path.join(GOLD, "spmv_small.npz"))
    for name in _cases(g):
        y = oracle.spmv((g[f"{name}/Ap"], g[f"{name}/Aj"], g[f"{name}/Ax"]), g[f"{name}/x"])
        want = g[f"{name}/y"]
        assert y.dtype == want.dtype
        # csr_matvec is a sequential mul+add per row, same as the oracle
        assert np.array_equal(y, want), name


def _sha(p, j, x):
    h = hashlib.sha256()
    h.update(np.ascontiguousarray(p, dtype=np.int32).tobytes())
    h.update(np.ascontiguousarray(j, dtype=np.int32).tobytes())
    h.update(np.ascontiguousarray(x).tobytes())
    return h.hexdigest()


def test_golden_checksums_grid_and_c1():
    with open(os.path.join(GOLD, "checksums.json")) as f:
        sums = json.load(f)["cases"]
    for n in helpers.GRID_SIZES:
        for d in helpers.GRID_DENSITIES:
            A, B = helpers.grid_case(n, d)
            got = helpers.oracle_spgemm(A, B, nthreads=0)
            want = sums[f"grid_n{n}_d{d:g}"]
            assert got[1].size == want["nnzC"]
            assert _sha(*got) == want["sha256"], f"grid n={n} d={d}"
    A = helpers.profiler_matrix(4096, 4096, 1e-3)
    want = sums["c1_4096_d1e-3"]
    assert A.nnz == want["nnzA"]
    assert oracle.num_products(A.indptr, A.indices, A.indptr) == want["products"]
    assert _sha(*helpers.oracle_spgemm(A, A)) == want["sha256"]


# ---- (c) live scipy ---------------------------------------------------------------------

@pytest.mark.parametrize("n,d", [(32, 0.5), (128, 0.3), (512, 0.1), (1024, 0.01)])
@pytest.mark.parametrize("dt", [np.float32, np.float64])
def test_oracle_vs_live_scipy(n, d, dt):
    A, B = helpers.grid_case(n, d, dtype=dt)
    C = (A @ B).tocsr()
    C.sort_indices()
    helpers.assert_csr_bitwise(helpers.oracle_spgemm(A, B), helpers.triple(C), f"n={n} d={d}")
    # threaded oracle == single-threaded oracle (rows are independent)
    helpers.assert_csr_bitwise(helpers.oracle_spgemm(A, B, nthreads=0), helpers.triple(C), "threaded")


def test_oracle_edge_cases():
    # cancellation keeps an explicit zero (structural nnz), like scipy and cuSPARSE
    A = sp.csr_matrix(np.array([[1.0, 1.0]], dtype=np.float32))
    B = sp.csr_matrix(np.array([[2.0], [-2.0]], dtype=np.float32))
    p, j, x = helpers.oracle_spgemm(A, B)
    assert p.tolist() == [0, 1] and j.tolist() == [0] and x.tolist() == [0.0]
    # empty operands / ragged rows / a 1 x 1 case
    Z = sp.csr_matrix((5, 7), dtype=np.float32)
    R = helpers.random_csr(7, 3, 0.5, 1)
    p, j, x = helpers.oracle_spgemm(Z, R)
    assert p.tolist() == [0] * 6 and j.size == 0
    p, j, x = helpers.oracle_spgemm(helpers.random_csr(4, 5, 0.5, 2), sp.csr_matrix((5, 9), dtype=np.float32))
    assert p.tolist() == [0] * 5
    A, B = helpers.skewed_pair(k=5000, n=20000)
    C = (A @ B).tocsr()
    C.sort_indices()
    helpers.assert_csr_bitwise(helpers.oracle_spgemm(A, B), helpers.triple(C), "skewed")


def test_zero_sums_structural_vs_scipy():
    """The one place SciPy and cuSPARSE differ: entries whose sum is exactly zero.  SciPy's second
    pass skips them (`if (sums[head] != 0)`), cuSPARSE keeps the structural pattern.  The oracle
    gives the structural form by default and SciPy's with drop_zeros=True."""
    A = sp.csr_matrix(np.array([[1.0, 1.0]], np.float32))
    B = sp.csr_matrix(np.array([[2.0, 1.0], [-2.0, 1.0]], np.float32))
    p, j, x = oracle.spgemm(helpers.triple(A), helpers.triple(B), 2)
    assert p.tolist() == [0, 2] and j.tolist() == [0, 1] and x.tolist() == [0.0, 2.0]
    p, j, x = oracle.spgemm(helpers.triple(A), helpers.triple(B), 2, drop_zeros=True)
    C = (A @ B).tocsr()
    assert p.tolist() == C.indptr.tolist() == [0, 1] and j.tolist() == C.indices.tolist() == [1] and x.tolist() == [2.0]
    # signed zeros, explicit zeros, small-integer values (many exact cancellations)
    rng = np.random.default_rng(5)
    for dt in (np.float32, np.float64):
        A = helpers.random_csr(120, 90, 0.3, 31, dtype=dt)
        B = helpers.random_csr(90, 150, 0.3, 32, dtype=dt)
        for M in (A, B):
            M.data[:] = rng.integers(-2, 3, M.nnz).astype(dt)
            M.data[rng.random(M.nnz) < 0.15] = dt(-0.0)
        full = oracle.spgemm(helpers.triple(A), helpers.triple(B), 150)
        C = (A @ B).tocsr()
        C.sort_indices()
        helpers.assert_csr_bitwise(oracle.drop_zero_entries(full), (C.indptr, C.indices, C.data), "drop_zeros vs scipy")
        assert full[1].size > C.nnz
        # structural count = csr_matmat_maxnnz, SciPy's own first pass
        ones = lambda M: sp.csr_matrix((np.ones(M.nnz), M.indices, M.indptr), shape=M.shape)  # noqa: E731
        assert full[1].size == (ones(A) @ ones(B)).nnz
