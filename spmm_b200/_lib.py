"""ctypes binding of libb200sp.so (include/b200sp.h).

This is the only place the package touches native code.  There is NO fallback: if the shared
library is missing the import fails loudly, and if no CUDA device is present every operator
raises (b200sp_create returns B200SP_ERR_CUDA).
"""
import ctypes
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
# B200SP_LIB: another build of the same ABI (A/B timing of two library versions inside one GPU call)
LIB_PATH = os.environ.get("B200SP_LIB") or os.path.join(_HERE, "libb200sp.so")

OK = 0
ERR_INVALID, ERR_CUDA, ERR_ALLOC, ERR_OVERFLOW, ERR_STATE, ERR_IO = -1, -2, -3, -4, -5, -6
R_32F, R_64F = 0, 1
ORDER_ROW, ORDER_COL = 0, 1
ALG_REUSE_ESTIMATE = 0x100
MAX_BINS = 16

c_i64 = ctypes.c_int64
c_vp = ctypes.c_void_p
i32p = ctypes.POINTER(ctypes.c_int32)
i64p = ctypes.POINTER(ctypes.c_int64)


class Stats(ctypes.Structure):
    _fields_ = [
        ("num_products", c_i64), ("nnz_c", c_i64), ("workspace_bytes", c_i64),
        ("peak_workspace_bytes", c_i64), ("max_row_products", c_i64),
        ("num_chunks", ctypes.c_int32), ("num_bins", ctypes.c_int32),
        ("sym_bin_rows", c_i64 * MAX_BINS), ("num_bin_rows", c_i64 * MAX_BINS),
        ("kernel_launches", c_i64), ("ms_symbolic", ctypes.c_float), ("ms_numeric", ctypes.c_float),
    ]


# every symbol include/b200sp.h declares: (name, restype, argtypes)
SYMBOLS = [
    ("b200sp_version", ctypes.c_int, []),
    ("b200sp_create", ctypes.c_int, [ctypes.POINTER(c_vp), ctypes.c_int]),
    ("b200sp_destroy", ctypes.c_int, [c_vp]),
    ("b200sp_set_stream", ctypes.c_int, [c_vp, c_vp]),
    ("b200sp_set_allocator", ctypes.c_int, [c_vp, c_vp, c_vp, c_vp]),
    ("b200sp_release_workspace", ctypes.c_int, [c_vp]),
    ("b200sp_last_error", ctypes.c_char_p, [c_vp]),
    ("b200sp_status_string", ctypes.c_char_p, [ctypes.c_int]),
    ("b200sp_spgemm_symbolic", ctypes.c_int,
     [c_vp, c_i64, c_i64, c_i64, c_i64, c_i64, c_vp, c_vp, c_vp, c_vp, ctypes.c_int, ctypes.c_float,
      i64p, i64p, c_vp]),
    ("b200sp_spgemm_estimate", ctypes.c_int,
     [c_vp, ctypes.c_int, c_i64, c_i64, c_i64, c_i64, c_i64, c_vp, c_vp, c_vp, c_vp, ctypes.c_int, ctypes.c_float,
      c_vp, i64p, i64p, ctypes.POINTER(ctypes.c_int)]),
    ("b200sp_spgemm_single_pass", ctypes.c_int, [c_vp, ctypes.c_int, c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, i64p]),
    ("b200sp_spgemm_single_pass_direct", ctypes.c_int,
     [c_vp, ctypes.c_int, c_i64, c_i64, c_i64, c_i64, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp,
      c_i64, i64p, i64p]),
    ("b200sp_spgemm_numeric", ctypes.c_int, [c_vp, ctypes.c_int, c_vp, c_vp, c_vp, c_vp, c_vp]),
    ("b200sp_spgemm_host", ctypes.c_int,
     [c_vp, ctypes.c_int, c_i64, c_i64, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, ctypes.c_int,
      ctypes.c_float, c_vp, ctypes.POINTER(c_vp), ctypes.POINTER(c_vp), ctypes.POINTER(c_vp), i64p, i64p]),
    ("b200sp_free_host", ctypes.c_int, [c_vp]),
    ("b200sp_host_register", ctypes.c_int, [c_vp, ctypes.c_size_t]),
    ("b200sp_host_unregister", ctypes.c_int, [c_vp]),
    ("b200sp_memcpy_async", ctypes.c_int, [c_vp, c_vp, c_vp, ctypes.c_size_t, ctypes.c_int, c_vp]),
    ("b200sp_host_row_products", ctypes.c_int, [c_i64, c_i64, c_vp, c_vp, c_vp, c_vp, ctypes.c_int]),
    ("b200sp_spmv", ctypes.c_int, [c_vp, ctypes.c_int, c_i64, c_i64, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    ("b200sp_spmv_noalloc", ctypes.c_int, [c_vp, ctypes.c_int, c_i64, c_i64, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    ("b200sp_spmm", ctypes.c_int,
     [c_vp, ctypes.c_int, c_i64, c_i64, c_i64, c_i64, c_vp, c_vp, c_vp, c_vp, c_i64, c_vp, c_i64, c_vp, c_vp]),
    ("b200sp_spmv_host", ctypes.c_int,
     [c_vp, ctypes.c_int, c_i64, c_i64, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    ("b200sp_csr_is_canonical", ctypes.c_int, [c_vp, c_i64, c_i64, c_vp, c_vp, ctypes.POINTER(ctypes.c_int)]),
    ("b200sp_csr_sort_indices", ctypes.c_int, [c_vp, ctypes.c_int, c_i64, c_i64, c_vp, c_vp, c_vp]),
    ("b200sp_coo_to_csr", ctypes.c_int,
     [c_vp, ctypes.c_int, c_i64, c_i64, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, i64p]),
    ("b200sp_csr_sum_duplicates", ctypes.c_int,
     [c_vp, ctypes.c_int, c_i64, c_i64, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, i64p]),
    ("b200sp_csc_to_csr", ctypes.c_int, [c_vp, ctypes.c_int, c_i64, c_i64, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    ("b200sp_csr_to_csc", ctypes.c_int, [c_vp, ctypes.c_int, c_i64, c_i64, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    ("b200sp_csr_to_dense", ctypes.c_int, [c_vp, ctypes.c_int, c_i64, c_i64, c_vp, c_vp, c_vp, c_vp, ctypes.c_int]),
    ("b200sp_csr_max_abs_diff", ctypes.c_int,
     [c_vp, ctypes.c_int, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, ctypes.POINTER(ctypes.c_double)]),
    ("b200sp_get_stats", ctypes.c_int, [c_vp, ctypes.POINTER(Stats)]),
    ("b200sp_reset_stats", ctypes.c_int, [c_vp]),
    ("b200sp_read_csr_txt", ctypes.c_int,
     [ctypes.c_char_p, i64p, i64p, ctypes.POINTER(c_vp), ctypes.POINTER(c_vp), ctypes.POINTER(c_vp),
      ctypes.c_char_p, ctypes.c_size_t]),
    ("b200sp_write_csr_txt", ctypes.c_int,
     [ctypes.c_char_p, c_i64, c_i64, c_vp, c_vp, c_vp, ctypes.c_char_p, ctypes.c_size_t]),
]

_lib = None
_lock = threading.Lock()


def load():
    """Load libb200sp.so (once) and declare the prototypes.  Raises ImportError if it is absent."""
    global _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise ImportError(
                    f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
                    f"g.build()'` or `make -C spmm_b200/csrc`.  spmm_b200 has no CPU / library fallback.")
            lib = ctypes.CDLL(LIB_PATH)
            for name, res, args in SYMBOLS:
                fn = getattr(lib, name)  # AttributeError if the header and the library disagree
                fn.restype = res
                fn.argtypes = args
            _lib = lib
    return _lib


class B200spError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"b200sp status {status}: {message}")
        self.status = status


def check(handle, status):
    """Map a status code to the exception the reference's callers already catch
    (dense_vs_sparseGEMM/utils.py:160,168: OutOfMemoryError / MemoryError / RuntimeError)."""
    if status == OK:
        return
    lib = load()
    msg = (lib.b200sp_last_error(handle) or b"").decode() if handle else ""
    msg = msg or lib.b200sp_status_string(status).decode()
    if status == ERR_ALLOC:
        raise MemoryError(msg)
    if status == ERR_INVALID:
        raise ValueError(msg)
    if status == ERR_OVERFLOW:
        raise OverflowError(msg)
    raise B200spError(status, msg)


_tls = threading.local()


def handle(device):
    """One library handle per (host thread, device)."""
    lib = load()
    cache = getattr(_tls, "handles", None)
    if cache is None:
        cache = _tls.handles = {}
    h = cache.get(device)
    if h is None:
        hv = c_vp()
        st = lib.b200sp_create(ctypes.byref(hv), int(device))
        if st != OK:
            raise B200spError(st, "b200sp_create failed: no usable CUDA device (there is no CPU fallback)")
        h = cache[device] = hv
    return h


def new_handle(device):
    """A private library handle (own plan + workspace): the row-panel driver keeps one per
    sub-panel so that the symbolic plans of all panels stay alive until their numeric pass."""
    hv = c_vp()
    st = load().b200sp_create(ctypes.byref(hv), int(device))
    if st != OK:
        raise B200spError(st, "b200sp_create failed: no usable CUDA device (there is no CPU fallback)")
    return hv


def destroy_handle(h):
    load().b200sp_destroy(h)


def stats(device, h=None):
    s = Stats()
    h = h or handle(device)
    check(h, load().b200sp_get_stats(h, ctypes.byref(s)))
    return s
