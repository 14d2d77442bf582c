"""GPU parity tests of SpMV / SpMM against the oracle (sequential per-row sums).

Tolerance: the kernels reduce each row in a tree / by tiles, so values differ from the
sequential CPU sum by rounding only: |got - want| <= 1e-5 * (sum_j |a_ij x_j|) for fp32 and
1e-13 for fp64 (tighter than the upstream decimal=6 check, test_cusparse.py:630-695).
"""
import os

import numpy as np
import pytest
import scipy.sparse as sp
import torch

import helpers
from oracle import oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def b200():
    import spmm_b200
    return spmm_b200


def check(y, A, x, want, dtype):
    bound = (abs(A) @ np.abs(x)).astype(np.float64) + 1e-30
    tol = 1e-5 if dtype == np.float32 else 1e-13
    err = np.abs(y.astype(np.float64) - want.astype(np.float64)) / bound
    assert err.max() <= tol, f"max scaled error {err.max():.3e} at row {int(err.argmax())}"


def matrices():
    rng = np.random.default_rng(5)
    out = {
        "short_rows": helpers.profiler_matrix(20000, 30000, 3e-4, seed=1),
        "c2_like": helpers.profiler_matrix(1 << 15, 1 << 15, 10.5 / (1 << 15), seed=2),
        "long_rows": helpers.profiler_matrix(300, 50000, 0.2, seed=3),
        "tiny": helpers.profiler_matrix(3, 4, 0.5, seed=4),
    }
    # skew: one row with 200 K entries, many empty rows, the rest short (R-MAT like)
    m, n = 5000, 250000
    rows = [np.full(200000, 17)]
    cols = [np.sort(rng.choice(n, 200000, replace=False))]
    for i in range(0, m, 3):
        if i == 17:
            continue
        L = int(rng.integers(1, 12))
        rows.append(np.full(L, i))
        cols.append(np.sort(rng.choice(n, L, replace=False)))
    rows, cols = np.concatenate(rows), np.concatenate(cols)
    S = sp.csr_matrix((rng.standard_normal(rows.size).astype(np.float32), (rows, cols)), shape=(m, n))
    S.sort_indices()
    out["skewed"] = S
    # rows that begin and end exactly on the SpMV tile boundaries (4096 fp32 / 2048 fp64 nonzeros),
    # with empty rows sitting on the boundaries and a row of 3 full tiles
    lens = [4096, 0, 2048, 2048, 0, 0, 12288, 1, 4095, 0, 8192, 0]
    rows = np.concatenate([np.full(L, i) for i, L in enumerate(lens)])
    cols = np.concatenate([np.sort(rng.choice(20000, L, replace=False)) for L in lens])
    T = sp.csr_matrix((rng.standard_normal(rows.size).astype(np.float32), (rows, cols)), shape=(len(lens), 20000))
    T.sort_indices()
    out["tile_aligned"] = T
    return out


MATS = None


def get(name):
    global MATS
    if MATS is None:
        MATS = matrices()
    return MATS[name]


@pytest.mark.parametrize("kernel", ["auto", "merge", "vector", "vector4", "merge+scalar", "vector+scalar"])
@pytest.mark.parametrize("name", ["short_rows", "c2_like", "long_rows", "tiny", "skewed", "tile_aligned"])
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_spmv(b200, name, dtype, kernel, monkeypatch):
    monkeypatch.setenv("B200SP_SPMV_KERNEL", kernel)
    A = get(name).astype(dtype)
    rng = np.random.default_rng(9)
    x = rng.standard_normal(A.shape[1]).astype(dtype)
    a = b200.csr_matrix(A)
    y = b200.cusparse.spmv(a, torch.from_numpy(x).cuda())
    assert y.shape == (A.shape[0],)
    check(y.cpu().numpy(), A, x, oracle.spmv(helpers.triple(A), x), dtype)
    # alpha / beta with a caller-supplied y (test_cusparse.py:676-695)
    y0 = rng.standard_normal(A.shape[0]).astype(dtype)
    yt = torch.from_numpy(y0.copy()).cuda()
    z = b200.cusparse.spmv(a, torch.from_numpy(x).cuda(), y=yt, alpha=0.5, beta=0.25)
    assert z is yt
    want = oracle.spmv(helpers.triple(A), x, y=y0, alpha=0.5, beta=0.25)
    bound = 0.5 * (abs(A) @ np.abs(x)) + 0.25 * np.abs(y0) + 1e-30
    tol = 1e-5 if dtype == np.float32 else 1e-13
    assert (np.abs(z.cpu().numpy().astype(np.float64) - want) / bound).max() <= tol


def test_spmv_unaligned_arrays(b200, monkeypatch):
    """Index / value arrays that do not start on a 16-byte boundary (views into larger buffers)
    must take the 4-byte-load kernels, not the 128-bit ones."""
    A = get("c2_like")
    x = np.random.default_rng(9).standard_normal(A.shape[1]).astype(np.float32)
    want = oracle.spmv(helpers.triple(A), x)
    dev = torch.device("cuda")
    for shift in (1, 2, 3):
        jbuf = torch.empty(A.nnz + 8, dtype=torch.int32, device=dev)
        xbuf = torch.empty(A.nnz + 8, dtype=torch.float32, device=dev)
        jbuf[shift:shift + A.nnz] = torch.from_numpy(A.indices).to(dev)
        xbuf[shift:shift + A.nnz] = torch.from_numpy(A.data).to(dev)
        a = b200.csr_matrix((xbuf[shift:shift + A.nnz], jbuf[shift:shift + A.nnz], torch.from_numpy(A.indptr).to(dev)),
                            shape=A.shape)
        assert a.indices.data_ptr() % 16 != 0
        for kernel in ("merge", "vector4"):
            monkeypatch.setenv("B200SP_SPMV_KERNEL", kernel)
            y = b200.cusparse.spmv(a, torch.from_numpy(x).to(dev))
            check(y.cpu().numpy(), A, x, want, np.float32)


def test_spmv_deterministic(b200):
    A = get("skewed")
    x = torch.from_numpy(np.random.default_rng(1).standard_normal(A.shape[1]).astype(np.float32)).cuda()
    a = b200.csr_matrix(A)
    y1 = b200.cusparse.spmv(a, x).clone()
    y2 = b200.cusparse.spmv(a, x)
    assert torch.equal(y1, y2)


def test_matmul_dispatch_n_by_1(b200):
    """A @ x with x of shape (n, 1): the SpMM-with-one-column path of the reference's profilers
    (SpGEMM_vs_SpMV/profiler.py:490-493); output shape (m, 1)."""
    A = get("c2_like")
    x = np.random.default_rng(44).standard_normal((A.shape[1], 1)).astype(np.float32)
    a = b200.csr_matrix(A)
    y = a @ torch.from_numpy(x).cuda()
    assert tuple(y.shape) == (A.shape[0], 1)
    check(y.cpu().numpy()[:, 0], A, x[:, 0], oracle.spmv(helpers.triple(A), x[:, 0]), np.float32)
    y1 = a @ torch.from_numpy(x[:, 0].copy()).cuda()
    assert tuple(y1.shape) == (A.shape[0],)


def test_spmm_columns(b200):
    """test_cusparse.py:739-798 (TestSpmm): dense operand with several columns, alpha / beta."""
    A = get("short_rows")
    rng = np.random.default_rng(3)
    X = rng.standard_normal((A.shape[1], 3)).astype(np.float32)
    a = b200.csr_matrix(A)
    Y = b200.cusparse.spmm(a, torch.from_numpy(X).cuda(), alpha=0.5)
    assert tuple(Y.shape) == (A.shape[0], 3)
    for c in range(3):
        check(Y.cpu().numpy()[:, c], A, X[:, c], oracle.spmv(helpers.triple(A), X[:, c], alpha=0.5), np.float32)


def test_spmv_errors(b200):
    a = b200.csr_matrix(get("tiny"))
    with pytest.raises(ValueError):
        b200.cusparse.spmv(a, torch.zeros(5, device="cuda"))
    with pytest.raises(TypeError):
        b200.cusparse.spmv(get("tiny"), torch.zeros(4, device="cuda"))
    e = b200.csr_matrix(sp.csr_matrix((3, 4), dtype=np.float32))
    assert b200.cusparse.spmv(e, torch.ones(4, device="cuda")).tolist() == [0.0, 0.0, 0.0]


def test_host_spmv(b200):
    A = get("c2_like")
    x = np.random.default_rng(2).standard_normal(A.shape[1]).astype(np.float32)
    y = b200.host.spmv_host(A, x, A.shape)
    check(y, A, x, oracle.spmv(helpers.triple(A), x), np.float32)
