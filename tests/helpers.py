"""Shared input generators and comparison helpers for the tests.

Generators restate the reference's recipes:
  * grid_case: cupy_cusparse/gen_and_save_alg1_txt.py:16-24 (random CSR, seeds s / s+1,
    sort_indices) with scipy/numpy RNG in place of CuPy's;
  * profiler_matrix: SpGEMM_vs_SpMV/profiler.py:296-302 (default_rng(seed), standard_normal
    values, eliminate_zeros) -- the C1/C3 inputs, B built with the same seed so B == A.
"""
import numpy as np
import scipy.sparse as sp

from oracle import oracle

GRID_SIZES = [32, 64, 128, 256, 512, 1024]       # cupy_cusparse/run_all_alg1.sh:5-7
GRID_DENSITIES = [0.01, 0.1, 0.3, 0.5]


def random_csr(m, n, density, seed, dtype=np.float32):
    rs = np.random.RandomState(seed)
    M = sp.random(m, n, density=density, format="csr", dtype=dtype, random_state=rs)
    M.sort_indices()
    return M


def grid_case(n, d, seed=123, dtype=np.float32):
    return random_csr(n, n, d, seed, dtype), random_csr(n, n, d, seed + 1, dtype)


def profiler_matrix(m, n, density, seed=42, dtype=np.float32):
    rng = np.random.default_rng(seed)
    M = sp.random(m, n, density=density, format="csr", dtype=dtype, random_state=rng,
                  data_rvs=rng.standard_normal)
    M.eliminate_zeros()
    M.sort_indices()
    return M


def triple(M):
    return M.indptr, M.indices, M.data


def oracle_spgemm(A, B, alpha=1.0, nthreads=0):
    return oracle.spgemm(triple(A), triple(B), B.shape[1], alpha=alpha, nthreads=nthreads)


def assert_csr_bitwise(got, want, what=""):
    gp, gj, gx = got
    wp, wj, wx = want
    assert gp.shape == wp.shape and np.array_equal(gp, wp), f"{what}: indptr differs"
    assert gj.shape == wj.shape, f"{what}: nnz differs ({gj.size} vs {wj.size})"
    if not np.array_equal(gj, wj):
        bad = int(np.flatnonzero(gj != wj)[0])
        row = int(np.searchsorted(wp, bad, side="right") - 1)
        raise AssertionError(f"{what}: indices differ first at {bad} (row {row}): got {gj[bad]} want {wj[bad]}")
    assert gx.dtype == wx.dtype, f"{what}: dtype {gx.dtype} vs {wx.dtype}"
    it = np.uint32 if gx.dtype == np.float32 else np.uint64
    if not np.array_equal(gx.view(it), wx.view(it)):
        bad = np.flatnonzero(gx.view(it) != wx.view(it))
        b0 = int(bad[0])
        row = int(np.searchsorted(wp, b0, side="right") - 1)
        raise AssertionError(f"{what}: {bad.size} values differ bitwise, first at {b0} (row {row}, col {wj[b0]}): "
                             f"got {gx[b0]!r} want {wx[b0]!r}")


def skewed_pair(seed=7, k=5000, n=300000, dtype=np.float32):
    """A with row lengths from 1 to 4000 against a wide, sparse B: rows of C from a handful of
    entries up to ~80 K, so every numeric/symbolic bin (group, warp, CTA hash, global) is hit."""
    rng = np.random.default_rng(seed)
    lens = [0, 1, 2, 3, 5, 10, 17, 40, 64, 100, 200, 400, 700, 1200, 2000, 4000] * 3
    rows, cols = [], []
    for i, L in enumerate(lens):
        c = np.sort(rng.choice(k, size=L, replace=False))
        rows.append(np.full(L, i))
        cols.append(c)
    rows, cols = np.concatenate(rows), np.concatenate(cols)
    A = sp.csr_matrix((rng.standard_normal(rows.size).astype(dtype), (rows, cols)), shape=(len(lens), k))
    B = sp.random(k, n, density=20.0 / n, format="csr", dtype=dtype, random_state=rng, data_rvs=rng.standard_normal)
    A.sort_indices()
    B.sort_indices()
    return A, B


def coo_with_pairs(m, n, k, seed, dtype=np.float32, dup_fraction=0.3):
    """Unsorted COO triple in which every position occurs at most twice, so that the sum of
    duplicates does not depend on their order (fp addition commutes) and SciPy's unstable
    in-row sort cannot change the bits."""
    rng = np.random.default_rng(seed)
    k = min(k, m * n)
    lin = rng.choice(m * n, size=k, replace=False)
    dup = lin[rng.random(k) < dup_fraction]
    lin = np.concatenate([lin, dup])
    rng.shuffle(lin)
    row, col = (lin // n).astype(np.int32), (lin % n).astype(np.int32)
    return row, col, rng.standard_normal(lin.size).astype(dtype)


def coo_heavy_duplicates(m, n, k, seed, dtype=np.float32):
    """COO with many-fold duplicates (k draws with replacement, a few hot positions)."""
    rng = np.random.default_rng(seed)
    row = rng.integers(0, m, k).astype(np.int32)
    col = rng.integers(0, n, k).astype(np.int32)
    hot = rng.random(k) < 0.05
    row[hot], col[hot] = m // 2, n // 3
    return row, col, rng.standard_normal(k).astype(dtype)


def scipy_canonical_from_coo(row, col, val, shape):
    M = sp.coo_matrix((val, (row, col)), shape=shape).tocsr()
    M.sum_duplicates()
    M.sort_indices()
    return M


def hub_pair(seed=3, k=3000, n=200000, hubs=40, hub_len=6000, tail_len=12, hot_cols=4, dtype=np.float32):
    """Power-law shaped pair for the column-blocked heavy-row kernel (n too wide for one
    shared-memory accumulator): `hubs` rows of B with `hub_len` entries (long segments inside a
    column block: the piece-per-barrier path) and the rest with `tail_len` entries (a handful per
    block: the ordered window path); every B row also holds columns 0 .. hot_cols-1, so the same
    accumulator is hit by every segment (the order of the sums is what the test is about).
    A: heavy rows that mix hub and tail rows of B, plus a few rows made of tail rows only."""
    rng = np.random.default_rng(seed)
    rows, cols = [], []
    for i in range(k):
        L = hub_len if i < hubs else tail_len
        c = np.unique(np.concatenate([np.arange(hot_cols), rng.integers(hot_cols, n, size=L)]))
        rows.append(np.full(c.size, i))
        cols.append(c)
    rows, cols = np.concatenate(rows), np.concatenate(cols)
    B = sp.csr_matrix((rng.standard_normal(rows.size).astype(dtype), (rows, cols)), shape=(k, n))
    a_rows, a_cols = [], []
    shapes = [(30, 300), (40, 0), (0, 2500), (5, 1200), (1, 700), (12, 2900), (0, 600), (40, 2960)]
    for i, (nh, nt) in enumerate(shapes * 3):
        c = np.concatenate([rng.choice(hubs, size=nh, replace=False), hubs + rng.choice(k - hubs, size=nt, replace=False)])
        c = np.sort(c)
        a_rows.append(np.full(c.size, i))
        a_cols.append(c)
    a_rows, a_cols = np.concatenate(a_rows), np.concatenate(a_cols)
    A = sp.csr_matrix((rng.standard_normal(a_rows.size).astype(dtype), (a_rows, a_cols)), shape=(len(shapes) * 3, k))
    A.sort_indices()
    B.sort_indices()
    return A, B


def rmat_matrix(scale):
    """R-MAT recipe of BASELINE config C5: the generator bench.py uses."""
    import os
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if root not in sys.path:
        sys.path.insert(0, root)
    import bench
    return bench.rmat_matrix(scale)
