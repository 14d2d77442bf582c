"""ctypes front end of oracle/liboracle.so (spgemm_oracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product package (spmm_b200) never does.

The functions take and return numpy arrays / scipy-like (indptr, indices, data) triples and
restate the reference's CPU path `A @ B` / `A @ x` (SpGEMM_vs_SpMV/profiler.py:408,411) with
sorted rows; see spgemm_oracle.c for the algorithm and how it is pinned.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

_i32p = ctypes.POINTER(ctypes.c_int32)
_i64p = ctypes.POINTER(ctypes.c_int64)


def build(quiet=True):
    """Compile liboracle.so (and oracle/_ref when /root/reference exists)."""
    subprocess.run(["make", "-C", _HERE, "--no-print-directory"], check=True,
                   stdout=subprocess.DEVNULL if quiet else None,
                   stderr=subprocess.DEVNULL if quiet else None)


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        L = ctypes.CDLL(path)
        L.oracle_num_products.restype = ctypes.c_int64
        L.oracle_spgemm_symbolic.restype = ctypes.c_int64
        L.oracle_num_threads.restype = ctypes.c_int
        L.oracle_csr_is_canonical.restype = ctypes.c_int
        _LIB = L
    return _LIB


def _p(a, ty):
    return a.ctypes.data_as(ty)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def num_threads():
    return int(lib().oracle_num_threads())


def num_products(Ap, Aj, Bp):
    Ap, Aj, Bp = _i32(Ap), _i32(Aj), _i32(Bp)
    return int(lib().oracle_num_products(ctypes.c_int64(len(Ap) - 1), _p(Ap, _i32p),
                                         _p(Aj, _i32p), _p(Bp, _i32p)))


def row_products(Ap, Aj, Bp):
    Ap, Aj, Bp = _i32(Ap), _i32(Aj), _i32(Bp)
    out = np.zeros(len(Ap) - 1, dtype=np.int64)
    lib().oracle_row_products(ctypes.c_int64(len(Ap) - 1), _p(Ap, _i32p), _p(Aj, _i32p),
                              _p(Bp, _i32p), _p(out, _i64p))
    return out


def drop_zero_entries(C):
    """SciPy's csr_matmat emits an entry only `if (sums[head] != 0)` (sparsetools csr.h, second
    pass): results that are exactly +-0.0 - cancellation, products of explicit zeros - are not
    stored, although csr_matmat_maxnnz counted them.  cuSPARSE (the path the product replaces)
    keeps the structural pattern.  This filter turns the structural result into SciPy's."""
    Cp, Cj, Cx = C
    keep = Cx != 0
    rows = np.repeat(np.arange(len(Cp) - 1), np.diff(Cp))
    newp = np.zeros(len(Cp), dtype=Cp.dtype)
    np.cumsum(np.bincount(rows[keep], minlength=len(Cp) - 1), out=newp[1:])
    return newp, Cj[keep], Cx[keep]


def spgemm(A, B, n_cols, alpha=1.0, nthreads=1, drop_zeros=False):
    """C = alpha * (A @ B).  A, B are (indptr, indices, data) triples; n_cols = B.shape[1].

    Returns (indptr, indices, data) with sorted rows, dtype of A's data (float32 / float64).
    drop_zeros=False: structural pattern (cuSPARSE semantics, what the product implements);
    drop_zeros=True: SciPy's output, which omits entries whose sum is exactly zero.
    """
    Ap, Aj, Ax = A
    Bp, Bj, Bx = B
    Ap, Aj, Bp, Bj = _i32(Ap), _i32(Aj), _i32(Bp), _i32(Bj)
    dt = np.result_type(Ax.dtype, Bx.dtype)
    if dt not in (np.float32, np.float64):
        raise TypeError("oracle handles float32 / float64 only")
    Ax = np.ascontiguousarray(Ax, dtype=dt)
    Bx = np.ascontiguousarray(Bx, dtype=dt)
    m = len(Ap) - 1
    Cp = np.zeros(m + 1, dtype=np.int32)
    nnz = lib().oracle_spgemm_symbolic(ctypes.c_int64(m), ctypes.c_int64(n_cols), _p(Ap, _i32p),
                                       _p(Aj, _i32p), _p(Bp, _i32p), _p(Bj, _i32p),
                                       _p(Cp, _i32p))
    if nnz < 0:
        raise OverflowError("nnz(C) does not fit int32 indptr" if nnz == -1 else "oracle OOM")
    Cj = np.empty(nnz, dtype=np.int32)
    Cx = np.empty(nnz, dtype=dt)
    if dt == np.float32:
        fn, cty, vp = lib().oracle_spgemm_numeric_f32, ctypes.c_float, ctypes.POINTER(ctypes.c_float)
    else:
        fn, cty, vp = lib().oracle_spgemm_numeric_f64, ctypes.c_double, ctypes.POINTER(ctypes.c_double)
    fn.restype = ctypes.c_int
    rc = fn(ctypes.c_int64(m), ctypes.c_int64(n_cols), _p(Ap, _i32p), _p(Aj, _i32p), _p(Ax, vp),
            _p(Bp, _i32p), _p(Bj, _i32p), _p(Bx, vp), cty(alpha), _p(Cp, _i32p), _p(Cj, _i32p),
            _p(Cx, vp), ctypes.c_int(nthreads))
    if rc != 0:
        raise RuntimeError(f"oracle numeric failed ({rc})")
    return drop_zero_entries((Cp, Cj, Cx)) if drop_zeros else (Cp, Cj, Cx)


def spmv(A, x, y=None, alpha=1.0, beta=0.0):
    """alpha * (A @ x) + beta * y for 1-D x; sequential per-row sums (csr_matvec order)."""
    Ap, Aj, Ax = A
    Ap, Aj = _i32(Ap), _i32(Aj)
    dt = np.result_type(Ax.dtype, x.dtype)
    Ax = np.ascontiguousarray(Ax, dtype=dt)
    x = np.ascontiguousarray(x, dtype=dt).reshape(-1)
    m = len(Ap) - 1
    out = np.zeros(m, dtype=dt) if y is None else np.array(y, dtype=dt).reshape(-1).copy()
    if dt == np.float32:
        fn, cty, vp = lib().oracle_spmv_f32, ctypes.c_float, ctypes.POINTER(ctypes.c_float)
    else:
        fn, cty, vp = lib().oracle_spmv_f64, ctypes.c_double, ctypes.POINTER(ctypes.c_double)
    fn.restype = None
    fn(ctypes.c_int64(m), _p(Ap, _i32p), _p(Aj, _i32p), _p(Ax, vp), _p(x, vp), _p(out, vp),
       cty(alpha), cty(beta))
    return out


def is_canonical(Ap, Aj, ncols):
    Ap, Aj = _i32(Ap), _i32(Aj)
    return bool(lib().oracle_csr_is_canonical(ctypes.c_int64(len(Ap) - 1), ctypes.c_int64(ncols),
                                              _p(Ap, _i32p), _p(Aj, _i32p)))


# ---- format conversions (convert_oracle.c) ------------------------------------------------

def _fp(dt):
    return (ctypes.c_float, ctypes.POINTER(ctypes.c_float), "f32") if dt == np.float32 else \
        (ctypes.c_double, ctypes.POINTER(ctypes.c_double), "f64")


def coo_to_csr(row, col, val, shape):
    """Canonical CSR of a COO triple: sorted by (row, col), duplicates summed in input order."""
    row, col = _i32(row), _i32(col)
    dt = np.dtype(val.dtype)
    val = np.ascontiguousarray(val, dtype=dt)
    m, n = shape
    _, vp, sfx = _fp(dt)
    Cp = np.zeros(m + 1, dtype=np.int32)
    Cj = np.empty(len(row), dtype=np.int32)
    Cx = np.empty(len(row), dtype=dt)
    fn = getattr(lib(), "oracle_coo_to_csr_" + sfx)
    fn.restype = ctypes.c_int64
    k = fn(ctypes.c_int64(m), ctypes.c_int64(n), ctypes.c_int64(len(row)), _p(row, _i32p), _p(col, _i32p),
           _p(val, vp), _p(Cp, _i32p), _p(Cj, _i32p), _p(Cx, vp))
    if k < 0:
        raise ValueError("coordinate out of range" if k == -1 else "oracle OOM")
    return Cp, Cj[:k].copy(), Cx[:k].copy()


def csr_sum_duplicates(A, shape):
    """Canonical form of a CSR with unsorted rows / duplicates (same summation rule)."""
    Ap, Aj, Ax = A
    row = np.repeat(np.arange(shape[0], dtype=np.int32), np.diff(Ap))
    return coo_to_csr(row, Aj, Ax, shape)


def transpose_compressed(ptr, idx, val, nminor):
    """(ptr, idx, val) compressed along the other axis (csr_tocsc restated): CSC -> CSR with
    nminor = number of rows, or CSR -> CSC with nminor = number of columns."""
    ptr, idx = _i32(ptr), _i32(idx)
    dt = np.dtype(val.dtype)
    val = np.ascontiguousarray(val, dtype=dt)
    _, vp, sfx = _fp(dt)
    optr = np.zeros(nminor + 1, dtype=np.int32)
    oidx = np.empty(len(idx), dtype=np.int32)
    oval = np.empty(len(idx), dtype=dt)
    fn = getattr(lib(), "oracle_transpose_" + sfx)
    fn.restype = ctypes.c_int
    rc = fn(ctypes.c_int64(len(ptr) - 1), ctypes.c_int64(nminor), _p(ptr, _i32p), _p(idx, _i32p), _p(val, vp),
            _p(optr, _i32p), _p(oidx, _i32p), _p(oval, vp))
    if rc:
        raise ValueError("index out of range")
    return optr, oidx, oval


def csr_max_abs_diff(A, B):
    Ap, Aj, Ax = A
    Bp, Bj, Bx = B
    dt = np.dtype(Ax.dtype)
    _, vp, sfx = _fp(dt)
    fn = getattr(lib(), "oracle_csr_max_abs_diff_" + sfx)
    fn.restype = ctypes.c_double
    Ap, Aj, Bp, Bj = _i32(Ap), _i32(Aj), _i32(Bp), _i32(Bj)
    Ax, Bx = np.ascontiguousarray(Ax, dtype=dt), np.ascontiguousarray(Bx, dtype=dt)
    return float(fn(ctypes.c_int64(len(Ap) - 1), _p(Ap, _i32p), _p(Aj, _i32p), _p(Ax, vp), _p(Bp, _i32p),
                    _p(Bj, _i32p), _p(Bx, vp)))
