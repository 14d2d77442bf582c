"""b200sp_spmv_noalloc (the allocation-free merge-path SpMV with its scratch in a file-scope device
array) against the existing entry point b200sp_spmv, through the C ABI, on two calls in a row.

Collected last on purpose (a failure here hides nothing after it).
Tolerance: the one tests/test_gpu_spmv.py uses - |got - want| <= 1e-5 (fp32) / 1e-13 (fp64) times
sum_j |a_ij x_j| against the default kernel choice of b200sp_spmv, and bit-for-bit against
b200sp_spmv forced onto the same merge-path kernels.
"""
import ctypes

import numpy as np
import pytest
import torch

import helpers

pytestmark = pytest.mark.gpu


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _scalar(v, dtype):
    return (ctypes.c_float if dtype == np.float32 else ctypes.c_double)(v)


def _spmv(fn, L, h, A, x, dtype, alpha=None, beta=None, y0=None):
    """Calls fn (b200sp_spmv or b200sp_spmv_noalloc) with device copies of A and x; returns y on the host."""
    dt = L.R_32F if dtype == np.float32 else L.R_64F
    td = torch.float32 if dtype == np.float32 else torch.float64
    ap = torch.from_numpy(A.indptr.astype(np.int32)).cuda()
    aj = torch.from_numpy(A.indices.astype(np.int32)).cuda()
    ax = torch.from_numpy(A.data.astype(dtype)).cuda()
    xd = torch.from_numpy(x.astype(dtype)).cuda()
    y = torch.from_numpy(y0.astype(dtype)).cuda() if y0 is not None else torch.full((A.shape[0],), float("nan"), dtype=td, device="cuda")
    torch.cuda.synchronize()
    a = _scalar(alpha, dtype) if alpha is not None else None
    b = _scalar(beta, dtype) if beta is not None else None
    st = fn(h, dt, A.shape[0], A.shape[1], A.nnz, _ptr(ap), _ptr(aj), _ptr(ax), _ptr(xd), _ptr(y),
            ctypes.byref(a) if a is not None else None, ctypes.byref(b) if b is not None else None)
    L.check(h, st)
    torch.cuda.synchronize()
    return y.cpu().numpy()


def _close(got, want, A, x, dtype):
    bound = (abs(A).astype(np.float64) @ np.abs(x.astype(np.float64))) + 1e-30
    tol = 1e-5 if dtype == np.float32 else 1e-13
    err = np.abs(got.astype(np.float64) - want.astype(np.float64)) / bound
    assert err.max() <= tol, f"max scaled error {err.max():.3e} at row {int(err.argmax())}"


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_noalloc_matches_spmv_two_calls_in_a_row(dtype, monkeypatch):
    from spmm_b200 import _lib as L
    lib = L.load()
    h = L.new_handle(0)
    try:
        rng = np.random.default_rng(11)
        # call 1: many tiles, rows that span tiles (long rows) and short rows; call 2: a smaller,
        # differently shaped matrix, so that every scratch word of call 2 sits on call 1's leftovers
        mats = [helpers.profiler_matrix(300, 50000, 0.2, seed=3).astype(dtype),
                helpers.profiler_matrix(20000, 30000, 3e-4, seed=1).astype(dtype)]
        for A in mats:
            x = rng.standard_normal(A.shape[1]).astype(dtype)
            monkeypatch.delenv("B200SP_SPMV_KERNEL", raising=False)
            want = _spmv(lib.b200sp_spmv, L, h, A, x, dtype)
            got = _spmv(lib.b200sp_spmv_noalloc, L, h, A, x, dtype)
            _close(got, want, A, x, dtype)
            monkeypatch.setenv("B200SP_SPMV_KERNEL", "merge")
            want_merge = _spmv(lib.b200sp_spmv, L, h, A, x, dtype)
            assert np.array_equal(got.view(np.uint8), want_merge.view(np.uint8)), "not bit-identical to the merge-path entry"
            # alpha / beta form on the same scratch again
            y0 = rng.standard_normal(A.shape[0]).astype(dtype)
            want2 = _spmv(lib.b200sp_spmv, L, h, A, x, dtype, alpha=0.5, beta=-2.0, y0=y0)
            got2 = _spmv(lib.b200sp_spmv_noalloc, L, h, A, x, dtype, alpha=0.5, beta=-2.0, y0=y0)
            assert np.array_equal(got2.view(np.uint8), want2.view(np.uint8))
    finally:
        lib.b200sp_destroy(h)


def test_noalloc_allocates_nothing_and_rejects_oversize():
    from spmm_b200 import _lib as L
    lib = L.load()
    h = L.new_handle(0)
    try:
        A = helpers.profiler_matrix(4096, 4096, 2e-3, seed=5).astype(np.float32)
        x = np.ones(A.shape[1], np.float32)
        _spmv(lib.b200sp_spmv_noalloc, L, h, A, x, np.float32)  # resolves the symbol address
        s = L.Stats()
        lib.b200sp_get_stats(h, ctypes.byref(s))
        assert s.workspace_bytes == 0 and s.peak_workspace_bytes == 0, "the handle's arena must stay untouched"
        free0 = torch.cuda.mem_get_info()[0]
        ap = torch.from_numpy(A.indptr.astype(np.int32)).cuda()
        aj = torch.from_numpy(A.indices.astype(np.int32)).cuda()
        ax = torch.from_numpy(A.data).cuda()
        xd = torch.from_numpy(x).cuda()
        y = torch.empty(A.shape[0], dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        free1 = torch.cuda.mem_get_info()[0]
        for _ in range(3):
            L.check(h, lib.b200sp_spmv_noalloc(h, L.R_32F, A.shape[0], A.shape[1], A.nnz, _ptr(ap), _ptr(aj), _ptr(ax),
                                               _ptr(xd), _ptr(y), None, None))
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free1 <= free0
        # beyond the stated maximum (m + nnz > 65536 tiles of 2048 merge items): refused before any launch
        big_m = 65536 * 2048 + 1
        st = lib.b200sp_spmv_noalloc(h, L.R_32F, big_m, 8, 0, _ptr(ap), _ptr(aj), _ptr(ax), _ptr(xd), _ptr(y), None, None)
        assert st == L.ERR_INVALID
        assert b"exceeds the fixed scratch" in lib.b200sp_last_error(h)
        with pytest.raises(ValueError):
            L.check(h, st)
    finally:
        lib.b200sp_destroy(h)
