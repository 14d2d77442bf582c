"""Host-buffer entry points over the C ABI (numpy in, numpy out).

These are the calls the end-to-end numbers are measured through: host CSR arrays in, H2D,
kernels, D2H, host CSR arrays out -- the body of the reference's C++ drivers
(cupy_cusparse/spgemm_from_txt_alg1.cu:130-195) and of the profilers' timed closures
(SpGEMM_alg_comparison/profiler.py:210-213, SpGEMM_vs_SpMV/profiler.py:485-493).
"""
import ctypes

import numpy as np

from . import _lib

_vp = ctypes.c_void_p


def _ptr(a):
    return _vp(a.ctypes.data) if a is not None and a.size else _vp(a.ctypes.data if a is not None else 0)


def _csr_arrays(M, dt=None):
    """(indptr, indices, data) C-contiguous int32/int32/float arrays from a scipy CSR or a triple."""
    if hasattr(M, "indptr"):
        p, j, x = M.indptr, M.indices, M.data
    else:
        p, j, x = M
    dt = np.dtype(dt or x.dtype)
    if dt not in (np.float32, np.float64):
        raise ValueError(f"Only float32 and float64 are supported (got {dt})")
    return (np.ascontiguousarray(p, dtype=np.int32), np.ascontiguousarray(j, dtype=np.int32),
            np.ascontiguousarray(x, dtype=dt))


def spgemm_host(A, B, shape, alg=0, chunk_fraction=0.2, alpha=1.0, device=0, copy=True):
    """C = alpha * A @ B through b200sp_spgemm_host.

    A, B: scipy CSR matrices or (indptr, indices, data) triples; shape = (m, k, n).
    Passing the same object for A and B uploads it once.  Returns (indptr, indices, data,
    num_products).  With copy=False the arrays are views of the handle's pinned staging
    buffers, valid until the next *_host call.
    """
    m, k, n = (int(s) for s in shape)
    Ap, Aj, Ax = _csr_arrays(A)
    if B is A:
        Bp, Bj, Bx = Ap, Aj, Ax
    else:
        Bp, Bj, Bx = _csr_arrays(B, Ax.dtype)
    dt = Ax.dtype
    lib = _lib.load()
    h = _lib.handle(device)
    _lib.check(h, lib.b200sp_set_stream(h, None))
    cp, cj, cx = _vp(), _vp(), _vp()
    nnz, prods = ctypes.c_int64(0), ctypes.c_int64(0)
    abuf = np.array(alpha, dtype=dt)
    _lib.check(h, lib.b200sp_spgemm_host(
        h, _lib.R_32F if dt == np.float32 else _lib.R_64F, m, k, n, _ptr(Ap), _ptr(Aj), _ptr(Ax), _ptr(Bp),
        _ptr(Bj), _ptr(Bx), int(alg), float(chunk_fraction), _vp(abuf.ctypes.data), ctypes.byref(cp),
        ctypes.byref(cj), ctypes.byref(cx), ctypes.byref(nnz), ctypes.byref(prods)))
    nz = int(nnz.value)

    def view(ptr, count, ty):
        if count == 0:
            return np.empty(0, dtype=ty)
        buf = (ctypes.c_char * (count * np.dtype(ty).itemsize)).from_address(ptr.value)
        arr = np.frombuffer(buf, dtype=ty, count=count)
        return arr.copy() if copy else arr

    return view(cp, m + 1, np.int32), view(cj, nz, np.int32), view(cx, nz, dt), int(prods.value)


def spmv_host(A, x, shape, y=None, alpha=1.0, beta=0.0, device=0):
    """y = alpha * A @ x + beta * y through b200sp_spmv_host (1-D x)."""
    m, n = (int(s) for s in shape)
    Ap, Aj, Ax = _csr_arrays(A)
    dt = Ax.dtype
    x = np.ascontiguousarray(x, dtype=dt).reshape(-1)
    if x.size != n:
        raise ValueError("dimension mismatch")
    out = np.zeros(m, dtype=dt) if y is None else np.ascontiguousarray(y, dtype=dt).reshape(-1).copy()
    lib = _lib.load()
    h = _lib.handle(device)
    _lib.check(h, lib.b200sp_set_stream(h, None))
    a, b = np.array(alpha, dtype=dt), np.array(beta, dtype=dt)
    _lib.check(h, lib.b200sp_spmv_host(h, _lib.R_32F if dt == np.float32 else _lib.R_64F, m, n, Aj.size, _ptr(Ap),
                                       _ptr(Aj), _ptr(Ax), _ptr(x), _ptr(out), _vp(a.ctypes.data),
                                       _vp(b.ctypes.data)))
    return out


def read_csr_txt(prefix):
    """(indptr, indices, data) from <prefix>_{indptr,indices,data}.txt via the library's fast
    reader (no GPU needed).  Mirrors read_csr_txt_prefix (spgemm_from_txt_alg1.cu:52-70)."""
    lib = _lib.load()
    rows, nnz = ctypes.c_int64(0), ctypes.c_int64(0)
    hp, hj, hx = _vp(), _vp(), _vp()
    err = ctypes.create_string_buffer(512)
    st = lib.b200sp_read_csr_txt(str(prefix).encode(), ctypes.byref(rows), ctypes.byref(nnz), ctypes.byref(hp),
                                 ctypes.byref(hj), ctypes.byref(hx), err, 512)
    if st != _lib.OK:
        msg = err.value.decode() or lib.b200sp_status_string(st).decode()
        raise (OSError if st == _lib.ERR_IO else ValueError)(msg)
    try:
        def take(ptr, count, ty):
            if count == 0:
                return np.empty(0, dtype=ty)
            buf = (ctypes.c_char * (count * np.dtype(ty).itemsize)).from_address(ptr.value)
            return np.frombuffer(buf, dtype=ty, count=count).copy()
        return take(hp, rows.value + 1, np.int32), take(hj, nnz.value, np.int32), take(hx, nnz.value, np.float32)
    finally:
        for p in (hp, hj, hx):
            lib.b200sp_free_host(p)


def write_csr_txt(prefix, indptr, indices, data):
    """Write the reference's three-file txt-CSR ("%d" / "%.9g"; write_csr_txt_prefix, alg1.cu:72-76)."""
    lib = _lib.load()
    p = np.ascontiguousarray(indptr, dtype=np.int32)
    j = np.ascontiguousarray(indices, dtype=np.int32)
    x = np.ascontiguousarray(data, dtype=np.float32)
    err = ctypes.create_string_buffer(512)
    st = lib.b200sp_write_csr_txt(str(prefix).encode(), p.size - 1, j.size, _ptr(p), _ptr(j), _ptr(x), err, 512)
    if st != _lib.OK:
        raise OSError(err.value.decode() or lib.b200sp_status_string(st).decode())
