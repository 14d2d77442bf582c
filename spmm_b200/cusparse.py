"""Operator layer: the mirror of the reference's `cupyx.cusparse` functions on the hot path
(modify_src/cupy-src/cupyx/cusparse.py: spgemm :2007-2142, spmv :1373-1437, spmm :1440-1513),
same names, argument meaning and error behaviour, backed by libb200sp.so instead of cuSPARSE.
"""
import ctypes

import numpy as np
import torch

from . import _lib
from .sparse import csr_matrix, _TORCH_DT, _bind_stream


def check_availability(name):
    """`cupyx.cusparse.check_availability` (cusparse.py:120-143) for the operators built here."""
    return name in ("spgemm", "spmv", "spmm", "csrsort")


def _dt_code(np_dtype):
    return _lib.R_32F if np.dtype(np_dtype) == np.float32 else _lib.R_64F


def _cast_common_type(*xs):
    """numpy.promote_types over the operands (cusparse.py:52-56), restricted to f32/f64."""
    dts = [x.dtype if isinstance(x, csr_matrix) else np.dtype(str(x.dtype).replace("torch.", ""))
           for x in xs if x is not None]
    dt = np.result_type(*dts)
    if dt not in _TORCH_DT:
        raise ValueError(f"Only float32 and float64 are supported (got {dt})")
    out = []
    for x in xs:
        if x is None:
            out.append(None)
        elif isinstance(x, csr_matrix):
            out.append(x if x.dtype == dt else x.astype(dt))
        else:
            out.append(x if x.dtype == _TORCH_DT[dt] else x.to(_TORCH_DT[dt]))
    return out


def _scalar(value, np_dtype):
    """alpha / beta as a host pointer to one value (CUSPARSE_POINTER_MODE_HOST, cusparse.py:2048-2049)."""
    buf = np.array(value, dtype=np_dtype)
    return buf, buf.ctypes.data_as(ctypes.c_void_p)


class SpgemmPlan:
    """The two halves of `spgemm` apart: the structure pass has run (C.indptr and nnz(C) are
    known, the one host synchronisation of cusparse.py:2120-2127 is behind us), the value pass is
    `numeric()`.  The row-panel driver (multi.py) plans all its panels first, learns where each
    lands in the shared host buffer, then runs the value passes.  `handle=None` uses the calling
    thread's handle, which keeps ONE plan: pass a private handle (`_lib.new_handle`) to keep
    several plans alive."""

    def __init__(self, a, b, alg=0, chunk_fraction=0.2, handle=None, _estimate=None):
        m, k = a.shape
        _, n = b.shape
        self.a, self.b = _cast_common_type(a, b)
        self.shape = (m, n)
        self.dtype = self.a.dtype
        dev = self.device = self.a.device
        lib = _lib.load()
        self.h = handle if handle is not None else _lib.handle(dev.index)
        _bind_stream(self.h, dev)
        if _estimate is not None:  # spgemm() has just run b200sp_spgemm_estimate on these operands and this handle
            self.indptr = _estimate
            alg = int(alg) | _lib.ALG_REUSE_ESTIMATE
        else:
            self.indptr = torch.empty(m + 1, dtype=torch.int32, device=dev)  # cusparse.py:2039
        prods, c_nnz = ctypes.c_int64(0), ctypes.c_int64(0)
        _lib.check(self.h, lib.b200sp_spgemm_symbolic(
            self.h, m, k, n, self.a.nnz, self.b.nnz, self.a.indptr.data_ptr(), self.a.indices.data_ptr(),
            self.b.indptr.data_ptr(), self.b.indices.data_ptr(), int(alg), float(chunk_fraction),
            ctypes.byref(prods), ctypes.byref(c_nnz), self.indptr.data_ptr()))
        self.nnz = int(c_nnz.value)
        self.num_products = int(prods.value)

    def numeric(self, alpha=1, out=None):
        """Fill C.indices / C.data (allocated here from torch's allocator unless `out` =
        (indices, data) tensors with room for nnz entries is given) and return the csr_matrix."""
        dev, dt, nnz = self.device, self.dtype, self.nnz
        _bind_stream(self.h, dev)
        if out is None:
            try:
                c_indices = torch.empty(nnz, dtype=torch.int32, device=dev)  # cusparse.py:2128-2129
                c_data = torch.empty(nnz, dtype=_TORCH_DT[dt], device=dev)
            except torch.OutOfMemoryError as e:  # same except clause as the reference's callers
                raise MemoryError(str(e)) from e
        else:
            c_indices, c_data = out[0][:nnz], out[1][:nnz]
        _abuf, aptr = _scalar(alpha, dt)
        _lib.check(self.h, _lib.load().b200sp_spgemm_numeric(
            self.h, _dt_code(dt), aptr, self.a.data.data_ptr(), self.b.data.data_ptr(), c_indices.data_ptr(),
            c_data.data_ptr()))
        c = csr_matrix((c_data, c_indices, self.indptr), shape=self.shape)
        c._canonical = True
        return c


_DIRECT_MAX_BYTES = 4 << 30  # C allocated at m * n entries: only while that stays small


def _spgemm_direct(a, b, alpha):
    """b200sp_spgemm_single_pass_direct when the sizes alone say C will be well filled: expected
    products nnzA * nnzB / k >= m * n / 2, an n-column accumulator fits shared memory and m * n entries
    of C are affordable.  No host synchronisation before the kernel (the estimate's costs ~45 us of a
    1.25 ms product on C3).  Returns None when it does not apply."""
    import os
    m, k = a.shape
    _, n = b.shape
    if os.environ.get("B200SP_SINGLE_PASS") == "0" or os.environ.get("B200SP_SINGLE_PASS_DIRECT") == "0":
        return None
    a, b = _cast_common_type(a, b)
    s = a.data.element_size()
    if not (m > 0 and k > 0 and n * s <= 208 * 1024 and m * n <= 2 ** 31 - 1 and m * n * (4 + s) <= _DIRECT_MAX_BYTES
            and a.nnz * (b.nnz / k) >= 0.5 * m * n):
        return None
    dev, dt = a.device, a.dtype
    lib = _lib.load()
    h = _lib.handle(dev.index)
    _bind_stream(h, dev)
    cap = m * n
    try:
        indptr = torch.empty(m + 1, dtype=torch.int32, device=dev)
        c_indices = torch.empty(cap, dtype=torch.int32, device=dev)
        c_data = torch.empty(cap, dtype=_TORCH_DT[dt], device=dev)
    except torch.OutOfMemoryError as e:
        raise MemoryError(str(e)) from e
    _abuf, aptr = _scalar(alpha, dt)
    c_nnz, prods = ctypes.c_int64(0), ctypes.c_int64(0)
    _lib.check(h, lib.b200sp_spgemm_single_pass_direct(
        h, _dt_code(dt), m, k, n, a.nnz, b.nnz, a.indptr.data_ptr(), a.indices.data_ptr(), a.data.data_ptr(),
        b.indptr.data_ptr(), b.indices.data_ptr(), b.data.data_ptr(), aptr, indptr.data_ptr(), c_indices.data_ptr(),
        c_data.data_ptr(), cap, ctypes.byref(c_nnz), ctypes.byref(prods)))
    nnz = int(c_nnz.value)
    c = csr_matrix((c_data[:nnz], c_indices[:nnz], indptr), shape=(m, n))
    c._canonical = True
    return c


def _spgemm_single_pass(a, b, alpha):
    """b200sp_spgemm_estimate, then b200sp_spgemm_single_pass when the library chooses it.
    Returns the product, or (a, b, indptr) for the two-pass plan to continue from the estimate."""
    m, k = a.shape
    _, n = b.shape
    a, b = _cast_common_type(a, b)
    dev, dt = a.device, a.dtype
    lib = _lib.load()
    h = _lib.handle(dev.index)
    _bind_stream(h, dev)
    indptr = torch.empty(m + 1, dtype=torch.int32, device=dev)  # cusparse.py:2039
    prods, bound, single = ctypes.c_int64(0), ctypes.c_int64(0), ctypes.c_int(0)
    _lib.check(h, lib.b200sp_spgemm_estimate(
        h, _dt_code(dt), m, k, n, a.nnz, b.nnz, a.indptr.data_ptr(), a.indices.data_ptr(), b.indptr.data_ptr(),
        b.indices.data_ptr(), 0, 0.2, indptr.data_ptr(), ctypes.byref(prods), ctypes.byref(bound), ctypes.byref(single)))
    if not single.value:
        return a, b, indptr
    cap = int(bound.value)
    try:
        c_indices = torch.empty(cap, dtype=torch.int32, device=dev)
        c_data = torch.empty(cap, dtype=_TORCH_DT[dt], device=dev)
    except torch.OutOfMemoryError as e:  # same except clause as the reference's callers
        raise MemoryError(str(e)) from e
    _abuf, aptr = _scalar(alpha, dt)
    c_nnz = ctypes.c_int64(0)
    _lib.check(h, lib.b200sp_spgemm_single_pass(h, _dt_code(dt), aptr, a.data.data_ptr(), b.data.data_ptr(),
                                                c_indices.data_ptr(), c_data.data_ptr(), cap, ctypes.byref(c_nnz)))
    nnz = int(c_nnz.value)
    c = csr_matrix((c_data[:nnz], c_indices[:nnz], indptr), shape=(m, n))
    c._canonical = True
    return c


def spgemm(a, b, alpha=1, alg=0, chunk_fraction=0.2, verbose=False):
    """C = alpha * A * B for CSR matrices (reference: cusparse.py:2007-2142).

    alg: 3 -> chunked low-memory mode, 2 -> chunked, anything else -> default single pass
    (the mapping of cusparse.py:2052-2057).  chunk_fraction in (0, 1] bounds the share of the
    intermediate products handled per chunk.  Returns a csr_matrix whose arrays are torch
    tensors from the caller's (torch caching) allocator, sorted and duplicate free.
    """
    if not check_availability("spgemm"):
        raise RuntimeError("spgemm is not available.")
    if not isinstance(a, csr_matrix):
        raise TypeError("unsupported type (actual: {})".format(type(a)))
    if not isinstance(b, csr_matrix):
        raise TypeError("unsupported type (actual: {})".format(type(b)))
    assert a.ndim == b.ndim == 2
    assert a.has_canonical_format
    assert b.has_canonical_format
    if a.shape[1] != b.shape[0]:
        raise ValueError("mismatched shape")
    if alg not in (2, 3):
        alg = 0
    elif not (0.0 < float(chunk_fraction) <= 1.0):
        raise ValueError(f"chunk_fraction must be in (0,1], got {chunk_fraction}")
    elif verbose:
        print("USING ALG3 or ALG2")
    est = None
    if alg == 0:
        # default mode: products per row first; well-filled C takes the single-pass kernel (C.indices /
        # C.data allocated at the bound sum_i min(products_i, n), trimmed to views of nnz(C) entries -
        # the memory-for-time trade of cuSPARSE's ALG1); anything else continues with the two-pass plan
        c = _spgemm_direct(a, b, alpha)
        if c is not None:
            return c
        c = _spgemm_single_pass(a, b, alpha)
        if isinstance(c, csr_matrix):
            return c
        a, b, est = c
    plan = SpgemmPlan(a, b, alg=alg, chunk_fraction=chunk_fraction, _estimate=est)
    if verbose:
        s = _lib.stats(plan.device.index)
        print("workspace GB =", s.workspace_bytes / (1024 ** 3), "products =", plan.num_products)
    return plan.numeric(alpha)


def spmv(a, x, y=None, alpha=1, beta=0, transa=False):
    """y = alpha * A x + beta * y (reference: cusparse.py:1373-1437)."""
    if not check_availability("spmv"):
        raise RuntimeError("spmv is not available.")
    if not isinstance(a, csr_matrix):
        raise TypeError("unsupported type (actual: {})".format(type(a)))
    if transa:
        raise NotImplementedError("transa=True is outside the hot path (never used by the reference's profilers)")
    if a.shape[1] != len(x):
        raise ValueError("dimension mismatch")
    assert a.has_canonical_format
    m, n = a.shape
    y_in = y
    a, x, y = _cast_common_type(a, x, y)
    dev = a.device
    if y is None:
        y = torch.zeros(m, dtype=a.data.dtype, device=dev)  # cusparse.py:1413
    elif len(y) != m:
        raise ValueError("dimension mismatch")
    if a.nnz == 0:
        y.fill_(0)  # cusparse.py:1416-1418
        return y
    x = x.contiguous()
    h = _lib.handle(dev.index)
    _bind_stream(h, dev)
    _ab, ap = _scalar(alpha, a.dtype)
    _bb, bp = _scalar(beta, a.dtype)
    _lib.check(h, _lib.load().b200sp_spmv(h, _dt_code(a.dtype), m, n, a.nnz, a.indptr.data_ptr(), a.indices.data_ptr(),
                                          a.data.data_ptr(), x.data_ptr(), y.data_ptr(), ap, bp))
    if y_in is not None and y is not y_in:
        y_in.copy_(y)
        return y_in
    return y


def spmm(a, b, c=None, alpha=1, beta=0, transa=False, transb=False):
    """C = alpha * A B + beta * C with dense column-major B, C (reference: cusparse.py:1440-1513).
    With an (n, 1) operand this is the path the reference's "SpMV" numbers take
    (_csr.py:217-225, SpGEMM_vs_SpMV/profiler.py:490-493).  Returns an (m, ncols) tensor whose
    memory is column-major (a transposed view of a contiguous (ncols, m) buffer)."""
    if not check_availability("spmm"):
        raise RuntimeError("spmm is not available.")
    if not isinstance(a, csr_matrix):
        raise TypeError("unsupported type (actual: {})".format(type(a)))
    if transa or transb:
        raise NotImplementedError("transposed operands are outside the hot path")
    assert a.ndim == b.ndim == 2
    if a.shape[1] != b.shape[0]:
        raise ValueError("dimension mismatch")
    assert a.has_canonical_format
    m, k = a.shape
    ncols = b.shape[1]
    a, b, c = _cast_common_type(a, b, c)
    dev = a.device
    bt = b.t().contiguous()  # (ncols, k) row-major == (k, ncols) column-major: asfortranarray
    if c is None:
        ct = torch.zeros((ncols, m), dtype=a.data.dtype, device=dev)
    else:
        if c.shape[0] != m or c.shape[1] != ncols:
            raise ValueError("dimension mismatch")
        ct = c.t().contiguous()
    if a.nnz == 0:
        ct.fill_(0)
    elif ncols:
        h = _lib.handle(dev.index)
        _bind_stream(h, dev)
        _ab, ap = _scalar(alpha, a.dtype)
        _bb, bp = _scalar(beta, a.dtype)
        _lib.check(h, _lib.load().b200sp_spmm(h, _dt_code(a.dtype), m, k, a.nnz, ncols, a.indptr.data_ptr(),
                                              a.indices.data_ptr(), a.data.data_ptr(), bt.data_ptr(), k,
                                              ct.data_ptr(), m, ap, bp))
    out = ct.t()
    if c is not None:
        c.copy_(out)
        return c
    return out
