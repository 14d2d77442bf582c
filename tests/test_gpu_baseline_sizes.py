"""GPU parity at the FULL sizes of the BASELINE.json configs (round-1 tests stopped short of
C2, C4 and C5 >= scale 18).  Everything goes through the C ABI; SpGEMM results are compared
bitwise (indptr, indices and values) with the oracle run on the box's host cores, and for the
R-MAT squares additionally with the committed partition-independent checksums
(tests/golden/rmat_checksums.json, produced by scipy / the oracle on the CPU).

  C2  SpMV 2^20 x 2^20, 1.1e7 nonzeros (bench.spmv_matrix), 1-D and (n, 1) x, all four kernels
  C4  8192^2 CSR @ CSR, d = 1e-5 .. 1e-1, independent A and B, uniform[0,1) values
      (dense_vs_sparseGEMM/utils.py:199-207, 238-240); 5.5e9 products and a fully dense C at the top
  C5  R-MAT scale 18 A @ A bitwise (2.9e9 products, 1.28e9 entries); scale 20 streamed in
      sub-panels (multi.spgemm_panels): checksums of the whole product + sampled sub-panels bitwise
"""
import json
import os

import numpy as np
import pytest
import scipy.sparse as sp
import torch

import helpers
from helpers import assert_csr_bitwise
from oracle import oracle

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
M64 = (1 << 64) - 1


@pytest.fixture(scope="module")
def b200():
    import spmm_b200
    return spmm_b200


def _bench():
    import bench
    return bench


# --------------------------------------------------------------------------- C2 ----

@pytest.fixture(scope="module")
def c2():
    A = _bench().spmv_matrix()
    x = np.random.default_rng(44).standard_normal((A.shape[1], 1)).astype(np.float32)  # seed + 2, profiler.py:305-307
    want = oracle.spmv(helpers.triple(A), x[:, 0])
    bound = (abs(A) @ np.abs(x[:, 0])).astype(np.float64) + 1e-30
    return A, x, want, bound


@pytest.mark.parametrize("kernel", ["auto", "vector", "vector4", "merge", "vector+scalar", "merge+scalar"])
def test_c2_spmv_full_size(b200, c2, kernel, monkeypatch):
    """|got - want| <= 1e-5 * sum_j |a_ij x_j| per row (fp32; the sum order differs from the
    sequential CPU loop, see tests/test_gpu_spmv.py)."""
    A, x, want, bound = c2
    assert A.shape == (1 << 20, 1 << 20) and A.nnz > 10_900_000
    monkeypatch.setenv("B200SP_SPMV_KERNEL", kernel)
    a = b200.csr_matrix(A)
    y1 = b200.cusparse.spmv(a, torch.from_numpy(x[:, 0].copy()).cuda())          # 1-D x -> spmv
    y2 = a @ torch.from_numpy(x).cuda()                                          # (n, 1) x -> spmm, one column
    assert tuple(y1.shape) == (A.shape[0],) and tuple(y2.shape) == (A.shape[0], 1)
    for y in (y1.cpu().numpy(), y2.cpu().numpy()[:, 0]):
        err = np.abs(y.astype(np.float64) - want.astype(np.float64)) / bound
        assert err.max() <= 1e-5, f"{kernel}: max scaled error {err.max():.3e} at row {int(err.argmax())}"
    # run-to-run determinism (deterministic/test_deterministic.sh criterion)
    assert torch.equal(y1, b200.cusparse.spmv(a, torch.from_numpy(x[:, 0].copy()).cuda()))


# --------------------------------------------------------------------------- C4 ----

def c4_pair(d, n=8192):
    """dense_vs_sparseGEMM/utils.py:199-207: np.random.seed(42), then A and B drawn one after the
    other with scipy.sparse.random (uniform[0,1) values)."""
    np.random.seed(42)
    A = sp.random(n, n, density=d, format="csr", dtype=np.float32)
    B = sp.random(n, n, density=d, format="csr", dtype=np.float32)
    A.sort_indices()
    B.sort_indices()
    return A, B


@pytest.mark.parametrize("d", [1e-5, 1e-4, 1e-3, 1e-2, 1e-1])
def test_c4_dense_vs_sparse_sizes_bitwise(b200, d):
    A, B = c4_pair(d)
    a, b = b200.csr_matrix(A), b200.csr_matrix(B)
    c = b200.cusparse.spgemm(a, b)
    want = oracle.spgemm(helpers.triple(A), helpers.triple(B), B.shape[1], nthreads=0)
    got = (c.indptr.cpu().numpy(), c.indices.cpu().numpy(), c.data.cpu().numpy())
    assert_csr_bitwise(got, want, f"C4 d={d:g}")
    if d == 1e-1:
        assert got[1].size == 8192 * 8192  # C is fully dense at the top of the sweep
    c2 = a @ b  # operator form, the call of utils.py:278
    assert torch.equal(c2.data, c.data) and torch.equal(c2.indices, c.indices)


# --------------------------------------------------------------------------- C5 ----

def golden(scale):
    with open(os.path.join(ROOT, "tests", "golden", "rmat_checksums.json")) as f:
        return json.load(f)[f"rmat{scale}"]


def device_checksums(r0, c):
    """(nnz, rowsum, colsum, colval) mod 2^64 of a device sub-panel whose first row is r0 - the
    definitions of tests/golden/make_rmat_checksums.py in torch int64 (two's complement wraps)."""
    cnt = (c.indptr[1:] - c.indptr[:-1]).to(torch.int64)
    rows = torch.arange(r0 + 1, r0 + 1 + cnt.numel(), dtype=torch.int64, device=cnt.device)
    rowsum = int((rows * cnt).sum().item()) & M64
    colsum, colval = 0, 0
    step = 1 << 28  # bounded temporaries
    for s in range(0, c.nnz, step):
        col1 = c.indices[s:s + step].to(torch.int64) + 1
        bits1 = (c.data[s:s + step].view(torch.int32).to(torch.int64) & 0xFFFFFFFF) + 1
        colsum = (colsum + int(col1.sum().item())) & M64
        colval = (colval + int((col1 * bits1).sum().item())) & M64
    return c.nnz, rowsum, colsum, colval


def test_c5_rmat18_bitwise_and_checksums(b200):
    A = helpers.rmat_matrix(18)
    g = golden(18)
    assert A.nnz == g["nnzA"]
    a = b200.csr_matrix(A)
    c = b200.cusparse.spgemm(a, a)
    chk = device_checksums(0, c)
    assert dict(zip(("nnz", "rowsum", "colsum", "colval"), chk)) == {k: g[k] for k in ("nnz", "rowsum", "colsum", "colval")}
    want = oracle.spgemm(helpers.triple(A), helpers.triple(A), A.shape[1], nthreads=0)
    got = (c.indptr.cpu().numpy(), c.indices.cpu().numpy(), c.data.cpu().numpy())
    assert_csr_bitwise(got, want, "R-MAT 18")


def test_c5_rmat20_streamed_checksums_and_sampled_panels(b200):
    from spmm_b200 import multi
    A = helpers.rmat_matrix(20)
    g = golden(20)
    assert A.nnz == g["nnzA"]
    m, n = A.shape
    b = b200.csr_matrix(A)
    T = helpers.triple(A)
    rp = multi.row_products(A.indptr, A.indices, A.indptr)
    tot = [0, 0, 0, 0]
    npanels = 0
    rng = np.random.default_rng(20)
    for p0, p1, c in multi.spgemm_panels(T, b, (m, n, n), row_prods=rp):
        part = device_checksums(p0, c)
        tot = [(t + v) & M64 for t, v in zip(tot, part)]
        # a sampled row range of every sub-panel, bitwise against the oracle
        r0 = p0 + int(rng.integers(0, max(1, p1 - p0 - 64)))
        r1 = min(p1, r0 + 64)
        P = A[r0:r1]
        want = oracle.spgemm(helpers.triple(P), T, n, nthreads=0)
        cp = c.indptr[r0 - p0:r1 - p0 + 1].cpu().numpy().astype(np.int64)
        lo, hi = int(cp[0]), int(cp[-1])
        got = ((cp - lo).astype(np.int32), c.indices[lo:hi].cpu().numpy(), c.data[lo:hi].cpu().numpy())
        assert_csr_bitwise(got, want, f"R-MAT 20 rows [{r0},{r1}) of sub-panel [{p0},{p1})")
        npanels += 1
        del c
    assert npanels >= 5  # nnz(C) = 9.7e9: at least five int32-sized sub-panels
    assert dict(zip(("nnz", "rowsum", "colsum", "colval"), tot)) == {k: g[k] for k in ("nnz", "rowsum", "colsum", "colval")}
