"""Device CSR container: the host-side mirror of `cupyx.scipy.sparse.csr_matrix` for the
SpGEMM / SpMV hot path (reference: modify_src/cupy-src/cupyx/scipy/sparse/_csr.py,
_compressed.py:194-300).  Arrays are torch CUDA tensors (torch is only the device-memory /
stream plumbing); all arithmetic goes through libb200sp.so.

Supported surface = what the reference's profilers and checkers touch: construction from a
SciPy matrix (H2D, `_compressed.py:225-231`), from `(data, indices, indptr)`, from a shape;
`.data/.indices/.indptr/.shape/.dtype/.nnz/.ndim`, `has_canonical_format`, `sum_duplicates()`,
`sort_indices()`, `get()`, `toarray()`, `@` / `*` with CSR (-> spgemm), 1-D (-> spmv) and 2-D
dense (-> spmm) operands (`_csr.py:151-229`).
"""
import ctypes

import numpy as np
import torch

from . import _lib

_TORCH_DT = {np.dtype(np.float32): torch.float32, np.dtype(np.float64): torch.float64}
_NP_DT = {v: k for k, v in _TORCH_DT.items()}


def _device():
    if not torch.cuda.is_available():
        raise RuntimeError("spmm_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def _check_dtype(dt):
    dt = np.dtype(dt)
    if dt not in _TORCH_DT:
        # the reference rejects everything but ?fdFD (_compressed.py:278-281); complex / bool are
        # outside the hot path and float16 is unreachable there too
        raise ValueError(f"Only float32 and float64 are supported (got {dt})")
    return dt


def _rtype(dt):
    return _lib.R_32F if np.dtype(dt) == np.float32 else _lib.R_64F


def _bind_stream(h, dev):
    _lib.check(h, _lib.load().b200sp_set_stream(h, ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)))


def _to_dev(arr, torch_dtype, dev):
    if isinstance(arr, torch.Tensor):
        return arr.to(device=dev, dtype=torch_dtype).contiguous()
    a = np.ascontiguousarray(arr)
    return torch.from_numpy(a).to(device=dev, dtype=torch_dtype, non_blocking=False)


class csr_matrix:
    format = "csr"
    ndim = 2

    def __init__(self, arg1, shape=None, dtype=None, copy=False):
        dev = _device()
        self._canonical = None
        if isinstance(arg1, _via_csr):  # csr_matrix(coo) / csr_matrix(csc): convert on the device
            arg1 = arg1.tocsr()
        if isinstance(arg1, csr_matrix):
            self.data = arg1.data.clone() if copy else arg1.data
            self.indices = arg1.indices.clone() if copy else arg1.indices
            self.indptr = arg1.indptr.clone() if copy else arg1.indptr
            self._shape = arg1.shape
            self._canonical = arg1._canonical
        elif hasattr(arg1, "tocsr") and hasattr(arg1, "nnz"):  # a scipy.sparse matrix
            m = arg1.tocsr()
            dt = _check_dtype(dtype or m.dtype)
            self.data = _to_dev(m.data.astype(dt, copy=False), _TORCH_DT[dt], dev)
            self.indices = _to_dev(m.indices.astype(np.int32, copy=False), torch.int32, dev)
            self.indptr = _to_dev(m.indptr.astype(np.int32, copy=False), torch.int32, dev)
            self._shape = tuple(int(s) for s in m.shape)
        elif isinstance(arg1, tuple) and len(arg1) == 2 and all(np.isscalar(s) for s in arg1):
            dt = _check_dtype(dtype or np.float32)  # csr_matrix((m, n)): empty, _compressed.py:215-223
            self._shape = (int(arg1[0]), int(arg1[1]))
            self.data = torch.empty(0, dtype=_TORCH_DT[dt], device=dev)
            self.indices = torch.empty(0, dtype=torch.int32, device=dev)
            self.indptr = torch.zeros(self._shape[0] + 1, dtype=torch.int32, device=dev)
            self._canonical = True
        elif isinstance(arg1, tuple) and len(arg1) == 3:
            data, indices, indptr = arg1
            if isinstance(data, torch.Tensor):
                dt = _check_dtype(dtype or _NP_DT.get(data.dtype, None) or np.float32)
            else:
                dt = _check_dtype(dtype or np.asarray(data).dtype)
            self.data = _to_dev(data, _TORCH_DT[dt], dev)
            self.indices = _to_dev(indices, torch.int32, dev)
            self.indptr = _to_dev(indptr, torch.int32, dev)
            if shape is None:
                ncols = int(self.indices.max().item()) + 1 if self.indices.numel() else 0
                shape = (self.indptr.numel() - 1, ncols)
            self._shape = (int(shape[0]), int(shape[1]))
            if self.indptr.numel() != self._shape[0] + 1:
                raise ValueError("index pointer size ({}) should be ({})".format(self.indptr.numel(), self._shape[0] + 1))
            if self.data.numel() != self.indices.numel():
                raise ValueError("indices and data should have the same size")
        else:
            raise ValueError("Unsupported initializer format")

    # -- basic attributes ---------------------------------------------------------------
    @property
    def shape(self):
        return self._shape

    @property
    def dtype(self):
        return _NP_DT[self.data.dtype]

    @property
    def nnz(self):
        return int(self.data.numel())

    @property
    def device(self):
        return self.data.device

    def _handle(self):
        h = _lib.handle(self.device.index)
        _bind_stream(h, self.device)
        return h

    @property
    def has_canonical_format(self):
        """Sorted, duplicate-free rows (the check kernel of _compressed.py:177-192); cached."""
        if self._canonical is None:
            res = ctypes.c_int(0)
            h = self._handle()
            _lib.check(h, _lib.load().b200sp_csr_is_canonical(
                h, self.shape[0], self.shape[1], self.indptr.data_ptr(), self.indices.data_ptr(), ctypes.byref(res)))
            self._canonical = bool(res.value)
        return self._canonical

    def sort_indices(self):
        """In-row column sort in place (reference: cupyx/cusparse.py:832-866 csrsort)."""
        if self.nnz:
            h = self._handle()
            dt = _lib.R_32F if self.dtype == np.float32 else _lib.R_64F
            _lib.check(h, _lib.load().b200sp_csr_sort_indices(
                h, dt, self.shape[0], self.nnz, self.indptr.data_ptr(), self.indices.data_ptr(), self.data.data_ptr()))
        self._canonical = None

    def sum_duplicates(self):
        """Sorted rows without duplicate entries, in place (_compressed.py:971-991).  No-op for
        canonical matrices (every matrix the reference's profilers build is); otherwise the rows
        are sorted and duplicates summed in storage order (b200sp_csr_sum_duplicates)."""
        if self.has_canonical_format:
            return
        m, n = self.shape
        cp = torch.empty(m + 1, dtype=torch.int32, device=self.device)
        cj = torch.empty(self.nnz, dtype=torch.int32, device=self.device)
        cx = torch.empty(self.nnz, dtype=self.data.dtype, device=self.device)
        out = ctypes.c_int64(0)
        h = self._handle()
        _lib.check(h, _lib.load().b200sp_csr_sum_duplicates(
            h, _rtype(self.dtype), m, n, self.nnz, self.indptr.data_ptr(), self.indices.data_ptr(),
            self.data.data_ptr(), cp.data_ptr(), cj.data_ptr(), cx.data_ptr(), ctypes.byref(out)))
        self.indptr, self.indices, self.data = cp, cj[:out.value].clone(), cx[:out.value].clone()
        self._canonical = True

    # -- transfers --------------------------------------------------------------------------
    def get(self):
        import scipy.sparse as sp
        return sp.csr_matrix((self.data.cpu().numpy(), self.indices.cpu().numpy(), self.indptr.cpu().numpy()),
                             shape=self.shape)

    def toarray(self, order=None):
        """Dense torch tensor with the same values (_csr.py:383-420; used by the numerical_error
        checks, numerical_error/error.py:27).  order 'C' (default) or 'F'."""
        order = "C" if order is None else order.upper()
        if order not in ("C", "F"):
            raise ValueError("order not understood")
        m, n = self.shape
        buf = torch.empty(m * n, dtype=self.data.dtype, device=self.device)
        if m * n:
            h = self._handle()
            _lib.check(h, _lib.load().b200sp_csr_to_dense(
                h, _rtype(self.dtype), m, n, self.indptr.data_ptr(), self.indices.data_ptr(), self.data.data_ptr(),
                buf.data_ptr(), _lib.ORDER_ROW if order == "C" else _lib.ORDER_COL))
        return buf.view(m, n) if order == "C" else buf.view(n, m).t()

    def _transposed_arrays(self):
        """(ptr, idx, val) of the CSC form = the CSR of the transpose (b200sp_csr_to_csc)."""
        m, n = self.shape
        ptr = torch.empty(n + 1, dtype=torch.int32, device=self.device)
        idx = torch.empty(self.nnz, dtype=torch.int32, device=self.device)
        val = torch.empty(self.nnz, dtype=self.data.dtype, device=self.device)
        h = self._handle()
        _lib.check(h, _lib.load().b200sp_csr_to_csc(
            h, _rtype(self.dtype), m, n, self.nnz, self.indptr.data_ptr(), self.indices.data_ptr(),
            self.data.data_ptr(), ptr.data_ptr(), idx.data_ptr(), val.data_ptr()))
        return ptr, idx, val

    def tocsc(self, copy=False):
        """_csr.py:455-477 (cusparse.csr2csc)."""
        ptr, idx, val = self._transposed_arrays()
        return csc_matrix((val, idx, ptr), shape=self.shape)

    def tocoo(self, copy=False):
        """_csr.py:433-453: row indices expanded from indptr."""
        counts = (self.indptr[1:] - self.indptr[:-1]).long()
        row = torch.repeat_interleave(torch.arange(self.shape[0], dtype=torch.int32, device=self.device), counts)
        return coo_matrix((self.data.clone() if copy else self.data, (row, self.indices)), shape=self.shape)

    def transpose(self, axes=None, copy=False):
        """_csr.py:512-540: the CSR of the transpose (the reference returns the CSC view of the
        same arrays; here the arrays are re-ordered so that the result can feed spgemm)."""
        if axes is not None:
            raise ValueError("swapping dimensions is the only logical permutation")
        ptr, idx, val = self._transposed_arrays()
        return csr_matrix((val, idx, ptr), shape=(self.shape[1], self.shape[0]))

    @property
    def T(self):
        return self.transpose()

    def tocsr(self, copy=False):
        return self.copy() if copy else self

    def copy(self):
        return csr_matrix(self, copy=True)

    def astype(self, dt):
        dt = _check_dtype(dt)
        out = csr_matrix((self.data.to(_TORCH_DT[dt]), self.indices, self.indptr), shape=self.shape)
        out._canonical = self._canonical
        return out

    # -- products (the hot path) -------------------------------------------------------------
    def __mul__(self, other):
        from . import cusparse
        if isinstance(other, csr_matrix):  # _csr.py:156-160
            self.sum_duplicates()
            other.sum_duplicates()
            return cusparse.spgemm(self, other)
        if np.isscalar(other):
            self.sum_duplicates()
            out = csr_matrix((self.data * other, self.indices, self.indptr), shape=self.shape)
            out._canonical = self._canonical
            return out
        if isinstance(other, (csc_matrix, coo_matrix)):  # _csr.py:161-164: convert, then spgemm
            return self * other.tocsr()
        if hasattr(other, "tocsr") and hasattr(other, "nnz"):  # scipy matrix: upload, then spgemm
            return self * csr_matrix(other)
        if isinstance(other, np.ndarray):
            other = torch.from_numpy(np.ascontiguousarray(other)).to(self.device)
        if isinstance(other, torch.Tensor):
            self.sum_duplicates()
            if other.ndim == 1:  # _csr.py:185-216
                return cusparse.spmv(self, other)
            if other.ndim == 2:  # _csr.py:217-225: column-major dense operand
                return cusparse.spmm(self, other)
            raise ValueError("could not interpret dimensions")
        return NotImplemented

    def __matmul__(self, other):  # _base.py:130-134
        if np.isscalar(other):
            raise ValueError("Scalar operands are not allowed, use '*' instead")
        return self.__mul__(other)

    dot = __mul__


class _via_csr:
    """Behaviour shared by the non-CSR containers: every product converts to CSR first, as the
    reference does on CUDA 12 (_csc.py:120-127 `self.tocsr()` -> spgemm; _base.py:105-134)."""
    ndim = 2

    @property
    def shape(self):
        return self._shape

    @property
    def dtype(self):
        return _NP_DT[self.data.dtype]

    @property
    def nnz(self):
        return int(self.data.numel())

    @property
    def device(self):
        return self.data.device

    def _handle(self):
        h = _lib.handle(self.device.index)
        _bind_stream(h, self.device)
        return h

    def __mul__(self, other):
        return self.tocsr().__mul__(other)

    def __matmul__(self, other):
        if np.isscalar(other):
            raise ValueError("Scalar operands are not allowed, use '*' instead")
        return self.tocsr().__mul__(other)

    dot = __mul__

    def toarray(self, order=None):
        return self.tocsr().toarray(order)

    def get(self):
        return self.tocsr().get().asformat(self.format)


class coo_matrix(_via_csr):
    """Device COO container (mirror of cupyx.scipy.sparse.coo_matrix, _coo.py): `.row/.col/.data`,
    built from a SciPy matrix or `(data, (row, col))`; `tocsr()` sorts by (row, col) and sums
    duplicates on the device (_coo.py:356-459, 517-540)."""
    format = "coo"

    def __init__(self, arg1, shape=None, dtype=None, copy=False):
        dev = _device()
        if hasattr(arg1, "tocoo") and hasattr(arg1, "nnz") and not isinstance(arg1, (csr_matrix, _via_csr)):
            c = arg1.tocoo()
            dt = _check_dtype(dtype or c.dtype)
            self.data = _to_dev(c.data.astype(dt, copy=False), _TORCH_DT[dt], dev)
            self.row = _to_dev(c.row.astype(np.int32, copy=False), torch.int32, dev)
            self.col = _to_dev(c.col.astype(np.int32, copy=False), torch.int32, dev)
            self._shape = tuple(int(v) for v in c.shape)
        elif isinstance(arg1, tuple) and len(arg1) == 2 and isinstance(arg1[1], tuple):
            data, (row, col) = arg1
            if isinstance(data, torch.Tensor):
                dt = _check_dtype(dtype or _NP_DT.get(data.dtype, None) or np.float32)
            else:
                dt = _check_dtype(dtype or np.asarray(data).dtype)
            self.data = _to_dev(data, _TORCH_DT[dt], dev)
            self.row = _to_dev(row, torch.int32, dev)
            self.col = _to_dev(col, torch.int32, dev)
            if not (self.data.numel() == self.row.numel() == self.col.numel()):
                raise ValueError("row, column, and data array must all be the same length")
            if shape is None:
                if self.row.numel() == 0:
                    raise ValueError("cannot infer dimensions from zero sized index arrays")
                shape = (int(self.row.max().item()) + 1, int(self.col.max().item()) + 1)
            self._shape = (int(shape[0]), int(shape[1]))
        else:
            raise ValueError("Unsupported initializer format")

    def tocsr(self, copy=False):
        m, n = self.shape
        nnz = self.nnz
        cp = torch.empty(m + 1, dtype=torch.int32, device=self.device)
        cj = torch.empty(nnz, dtype=torch.int32, device=self.device)
        cx = torch.empty(nnz, dtype=self.data.dtype, device=self.device)
        out = ctypes.c_int64(0)
        h = self._handle()
        _lib.check(h, _lib.load().b200sp_coo_to_csr(
            h, _rtype(self.dtype), m, n, nnz, self.row.data_ptr(), self.col.data_ptr(), self.data.data_ptr(),
            cp.data_ptr(), cj.data_ptr(), cx.data_ptr(), ctypes.byref(out)))
        k = out.value
        res = csr_matrix((cx if k == nnz else cx[:k].clone(), cj if k == nnz else cj[:k].clone(), cp), shape=(m, n))
        res._canonical = True
        return res

    def tocsc(self, copy=False):
        return self.tocsr().tocsc()

    def tocoo(self, copy=False):
        return self

    def sum_duplicates(self):
        """In place: sorted by (row, col), duplicates summed (_coo.py:356-459)."""
        c = self.tocsr().tocoo()
        self.row, self.col, self.data = c.row, c.col, c.data

    def transpose(self, axes=None, copy=False):
        return coo_matrix((self.data, (self.col, self.row)), shape=(self.shape[1], self.shape[0]))

    @property
    def T(self):
        return self.transpose()


class csc_matrix(_via_csr):
    """Device CSC container (mirror of cupyx.scipy.sparse.csc_matrix, _csc.py): `.data/.indices`
    (row indices) `/.indptr` (column pointers); `tocsr()` re-orders on the device
    (_csc.py:304-326, cusparse.csc2csr)."""
    format = "csc"

    def __init__(self, arg1, shape=None, dtype=None, copy=False):
        dev = _device()
        if hasattr(arg1, "tocsc") and hasattr(arg1, "nnz") and not isinstance(arg1, (csr_matrix, _via_csr)):
            c = arg1.tocsc()
            dt = _check_dtype(dtype or c.dtype)
            self.data = _to_dev(c.data.astype(dt, copy=False), _TORCH_DT[dt], dev)
            self.indices = _to_dev(c.indices.astype(np.int32, copy=False), torch.int32, dev)
            self.indptr = _to_dev(c.indptr.astype(np.int32, copy=False), torch.int32, dev)
            self._shape = tuple(int(v) for v in c.shape)
        elif isinstance(arg1, tuple) and len(arg1) == 3:
            data, indices, indptr = arg1
            if isinstance(data, torch.Tensor):
                dt = _check_dtype(dtype or _NP_DT.get(data.dtype, None) or np.float32)
            else:
                dt = _check_dtype(dtype or np.asarray(data).dtype)
            self.data = _to_dev(data, _TORCH_DT[dt], dev)
            self.indices = _to_dev(indices, torch.int32, dev)
            self.indptr = _to_dev(indptr, torch.int32, dev)
            if shape is None:
                nrows = int(self.indices.max().item()) + 1 if self.indices.numel() else 0
                shape = (nrows, self.indptr.numel() - 1)
            self._shape = (int(shape[0]), int(shape[1]))
            if self.indptr.numel() != self._shape[1] + 1:
                raise ValueError("index pointer size ({}) should be ({})".format(self.indptr.numel(), self._shape[1] + 1))
            if self.data.numel() != self.indices.numel():
                raise ValueError("indices and data should have the same size")
        else:
            raise ValueError("Unsupported initializer format")

    def tocsr(self, copy=False):
        m, n = self.shape
        cp = torch.empty(m + 1, dtype=torch.int32, device=self.device)
        cj = torch.empty(self.nnz, dtype=torch.int32, device=self.device)
        cx = torch.empty(self.nnz, dtype=self.data.dtype, device=self.device)
        h = self._handle()
        _lib.check(h, _lib.load().b200sp_csc_to_csr(
            h, _rtype(self.dtype), m, n, self.nnz, self.indptr.data_ptr(), self.indices.data_ptr(),
            self.data.data_ptr(), cp.data_ptr(), cj.data_ptr(), cx.data_ptr()))
        return csr_matrix((cx, cj, cp), shape=(m, n))

    def tocsc(self, copy=False):
        return self

    def tocoo(self, copy=False):
        return self.tocsr().tocoo()

    def transpose(self, axes=None, copy=False):
        """_csc.py:337-365: the same arrays read as the CSR of the transpose."""
        return csr_matrix((self.data, self.indices, self.indptr), shape=(self.shape[1], self.shape[0]))

    @property
    def T(self):
        return self.transpose()


def max_abs_diff(a, b):
    """max |a - b| of two CSR matrices of the same shape without forming dense arrays: the fused
    form of `abs(C.toarray() - D.toarray()).max()` (numerical_error/error.py:25-29)."""
    if not (isinstance(a, csr_matrix) and isinstance(b, csr_matrix)):
        raise TypeError("max_abs_diff expects two csr_matrix operands")
    if a.shape != b.shape:
        raise ValueError("mismatched shape")
    if a.dtype != b.dtype:
        raise TypeError("mismatched dtype")
    a.sum_duplicates()
    b.sum_duplicates()
    res = ctypes.c_double(0.0)
    h = a._handle()
    _lib.check(h, _lib.load().b200sp_csr_max_abs_diff(
        h, _rtype(a.dtype), a.shape[0], a.indptr.data_ptr(), a.indices.data_ptr(), a.data.data_ptr(),
        b.indptr.data_ptr(), b.indices.data_ptr(), b.data.data_ptr(), ctypes.byref(res)))
    return res.value


def isspmatrix_csr(x):
    return isinstance(x, csr_matrix)


def isspmatrix(x):
    return isinstance(x, (csr_matrix, csc_matrix, coo_matrix))
