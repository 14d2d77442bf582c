"""`cupy` / `cupyx` alias layer: lets the reference's Python scripts run UNCHANGED on this library.

    python -m spmm_b200.compat /path/to/SpGEMM_alg_comparison/profiler.py --size 1024 --density 0.01 --runs 3
    (or, inside Python:  import spmm_b200.compat as c; c.install(); import cupy ...)

`install()` puts stand-ins for `cupy`, `cupy.cuda`, `cupy.cuda.runtime`, `cupy.cuda.memory`,
`cupy.cuda.profiler`, `cupy.random`, `cupyx`, `cupyx.scipy`, `cupyx.scipy.sparse`, `cupyx.cusparse`
and `cupyx.profiler` into `sys.modules`.  Only the surface the reference's scripts touch is there
(file:line of every use is cited beside the member); dense arrays are torch CUDA tensors (torch is
the device-memory plumbing of this repo), sparse products go to libb200sp through
spmm_b200.cusparse / spmm_b200.sparse.  Nothing here computes on the CPU except the random INPUT
generators, which the reference also runs through a library RNG (CuPy's; its streams cannot be
reproduced, so the inputs differ from a CuPy run while every script-level property - shapes,
densities, determinism under `cp.random.seed`, tables, exit codes - is the same).

Peak-VRAM accounting (SpGEMM_alg_comparison/profiler.py:82-143): `MemoryPool.used_bytes()` is the
growth of torch's allocated bytes plus the library's workspace since the pool was made;
`runtime.memGetInfo()` is the driver's (so the profilers' `free0 - min_free` sees torch's blocks and
the library's cudaMalloc'ed workspace alike); `free_all_blocks()` returns cached blocks to the
driver.  Run with PYTORCH_CUDA_ALLOC_CONF=expandable_segments:True (the launcher sets it) so that
torch requests memory in 2 MiB steps rather than 20 MiB blocks.
"""
import contextlib
import os
import runpy
import sys
import types

import numpy as np

_STATE = {"installed": False, "rng": None}


def _torch():
    import torch
    return torch


def _dev():
    torch = _torch()
    if not torch.cuda.is_available():
        raise RuntimeError("spmm_b200.compat needs a CUDA device (there is no CPU fallback)")
    return torch.device("cuda", torch.cuda.current_device())


def _np_dtype(dt):
    torch = _torch()
    if dt is None:
        return None
    if isinstance(dt, torch.dtype):
        return np.dtype(str(dt).replace("torch.", ""))
    return np.dtype(dt)


def _torch_dtype(dt):
    torch = _torch()
    if dt is None or isinstance(dt, torch.dtype):
        return dt
    return getattr(torch, np.dtype(dt).name)


def _make_ndarray_type():
    torch = _torch()

    class ndarray(torch.Tensor):
        """cupy.ndarray stand-in: a torch CUDA tensor with the few CuPy methods the scripts call."""

        def get(self):  # profiler.py:199-203 `cp.asarray(times).get()`
            return self.detach().cpu().numpy()

        def astype(self, dt, copy=True):  # gen_and_save_alg1_txt.py:9-11, numerical_error/error.py:14
            out = self.to(_torch_dtype(dt))
            return out.clone() if (copy and out is self) else out

        def toarray(self):
            return self

        @property
        def shape(self):  # printed in the result tables (profiler.py:229)
            return tuple(torch.Tensor.shape.__get__(self))

        def __str__(self):  # deterministic/cupy_alg1.py:33-36 writes str(C.indices) ... to the diffed file
            return np.array2string(self.get(), threshold=1000)

    return ndarray


def asarray(a, dtype=None):
    """cp.asarray (profiler.py:199; SpGEMM_vs_SpMV/profiler.py:492; dense_vs_sparseGEMM/utils.py:219)."""
    torch = _torch()
    nd = _STATE["ndarray"]
    if isinstance(a, torch.Tensor):
        t = a.to(_dev())
    else:
        t = torch.from_numpy(np.ascontiguousarray(np.asarray(a, dtype=_np_dtype(dtype)))).to(_dev())
    if dtype is not None:
        t = t.to(_torch_dtype(dtype))
    return t.as_subclass(nd)


def asnumpy(a):
    """cp.asnumpy (gen_and_save_alg1_txt.py:9-11)."""
    torch = _torch()
    return a.detach().cpu().numpy() if isinstance(a, torch.Tensor) else np.asarray(a)


# ------------------------------------------------------------------------- cupy.random ----

class RandomState:
    """cupy.random.RandomState stand-in on numpy's generator (host side; results are uploaded)."""

    def __init__(self, seed=None):
        self._rs = np.random.RandomState(seed)

    def standard_normal(self, size=None, dtype=float):  # SpGEMM_vs_SpMV/profiler.py:310-312
        return asarray(self._rs.standard_normal(size).astype(_np_dtype(dtype)))

    def uniform(self, low=0.0, high=1.0, size=None, dtype=float):
        return asarray(self._rs.uniform(low, high, size).astype(_np_dtype(dtype)))

    def rand(self, *size):
        return asarray(self._rs.rand(*size))

    def choice(self, a, size=None, replace=True):
        return asarray(self._rs.choice(a, size=size, replace=replace))


def _global_rs():
    if _STATE["rng"] is None:
        _STATE["rng"] = RandomState(None)
    return _STATE["rng"]


def _seed(seed=None):  # deterministic/cupy_alg1.py:15, numerical_error/error.py:10, range.py:24 (no argument)
    _STATE["rng"] = RandomState(seed)


# -------------------------------------------------------------------------- cupy.cuda ----

class _NullStream:
    def synchronize(self):  # profiler.py:104-105
        _torch().cuda.synchronize()


class Stream:
    null = _NullStream()


class Event:
    """cp.cuda.Event (deterministic/cupy_alg1.py:12,31-32)."""

    def __init__(self, *a, **k):
        self._e = _torch().cuda.Event(enable_timing=True)

    def record(self, stream=None):
        self._e.record()

    def synchronize(self):
        self._e.synchronize()


def _lib_workspace_bytes():
    from . import _lib
    try:
        return int(_lib.stats(_torch().cuda.current_device()).workspace_bytes)
    except Exception:
        return 0


class MemoryPool:
    """cp.cuda.MemoryPool as the profilers use it (profiler.py:97-131): a fresh accounting scope.
    Allocation itself stays with torch's caching allocator and the library's arena."""

    def __init__(self, *a, **k):
        torch = _torch()
        torch.cuda.synchronize()
        self._base_alloc = torch.cuda.memory_allocated()
        self._base_res = torch.cuda.memory_reserved()
        self._base_ws = _lib_workspace_bytes()

    def malloc(self, size):  # only ever passed to using_allocator
        raise NotImplementedError("allocation goes through torch's allocator")

    def used_bytes(self):
        return max(0, _torch().cuda.memory_allocated() - self._base_alloc) + max(0, _lib_workspace_bytes() - self._base_ws)

    def total_bytes(self):
        return max(0, _torch().cuda.memory_reserved() - self._base_res) + max(0, _lib_workspace_bytes() - self._base_ws)

    def free_all_blocks(self):
        """pool.free_all_blocks() (profiler.py:129): cached blocks and the library arena go back to the driver."""
        import gc
        from . import _lib
        gc.collect()
        torch = _torch()
        torch.cuda.synchronize()
        try:
            _lib.load().b200sp_release_workspace(_lib.handle(torch.cuda.current_device()))
        except Exception:
            pass
        torch.cuda.empty_cache()

    def n_free_blocks(self):
        return 0


class PinnedMemoryPool:
    def free_all_blocks(self):
        pass


@contextlib.contextmanager
def using_allocator(allocator=None):
    """cp.cuda.using_allocator (profiler.py:98): the scope starts from a clean allocator so that the
    profilers' `free0` does not include blocks cached by earlier runs."""
    MemoryPool().free_all_blocks()
    yield


class _Runtime:
    @staticmethod
    def memGetInfo():  # profiler.py:90,99
        return _torch().cuda.mem_get_info()

    @staticmethod
    def getDevice():  # profiler.py:182
        return _torch().cuda.current_device()

    @staticmethod
    def getDeviceProperties(dev):
        p = _torch().cuda.get_device_properties(dev)
        return {"name": p.name.encode(), "totalGlobalMem": p.total_memory, "multiProcessorCount": p.multi_processor_count,
                "major": p.major, "minor": p.minor}

    @staticmethod
    def deviceSynchronize():
        _torch().cuda.synchronize()


# ----------------------------------------------------------------- cupyx.scipy.sparse ----

def _wrap_sparse_classes():
    """csr / csc / coo containers whose arrays answer to the CuPy array methods and whose products
    come back as the same wrapped types."""
    from . import sparse as S
    nd = _STATE["ndarray"]

    def adopt(x):
        for name in ("data", "indices", "indptr", "row", "col"):
            v = getattr(x, name, None)
            if v is not None and not isinstance(v, nd):
                setattr(x, name, v.as_subclass(nd))
        return x

    def rewrap(out):
        torch = _torch()
        if isinstance(out, S.csr_matrix) and not isinstance(out, csr_matrix):
            w = csr_matrix.__new__(csr_matrix)
            w.__dict__.update(out.__dict__)
            return adopt(w)
        if isinstance(out, torch.Tensor) and not isinstance(out, nd):
            return out.as_subclass(nd)
        return out

    class _Mixin:
        def __matmul__(self, other):
            return rewrap(super().__matmul__(other))

        def __mul__(self, other):
            return rewrap(super().__mul__(other))

        dot = __mul__

        def toarray(self, order=None):
            return rewrap(super().toarray(order))

        def tocsr(self, copy=False):
            return rewrap(super().tocsr(copy))

        def asformat(self, fmt):
            return self if fmt == self.format else {"csr": csr_matrix, "csc": csc_matrix, "coo": coo_matrix}[fmt](self.get())

    class csr_matrix(_Mixin, S.csr_matrix):
        def __init__(self, *a, **k):
            super().__init__(*a, **k)
            adopt(self)

    class csc_matrix(_Mixin, S.csc_matrix):
        def __init__(self, *a, **k):
            super().__init__(*a, **k)
            adopt(self)

    class coo_matrix(_Mixin, S.coo_matrix):
        def __init__(self, *a, **k):
            super().__init__(*a, **k)
            adopt(self)

    return csr_matrix, csc_matrix, coo_matrix, rewrap


def _sparse_random(m, n, density=0.01, format="coo", dtype=None, random_state=None, data_rvs=None):
    """cupyx.scipy.sparse.random (_construct.py:321-374): k = int(density m n) distinct positions drawn
    without replacement, values from data_rvs(k) (default: uniform [0, 1)), returned in `format`.
    The draw is made on the host with numpy's RandomState (the scope's seeded global one when
    random_state is None, as cupy.random.seed fixes CuPy's) and uploaded once."""
    import scipy.sparse as sp
    if density < 0 or density > 1:
        raise ValueError("density expected to be 0 <= density <= 1")
    dt = _np_dtype(dtype) or np.dtype(np.float64)
    rs = random_state if isinstance(random_state, RandomState) else (RandomState(random_state) if random_state is not None else _global_rs())
    mn = m * n
    k = int(density * m * n)
    ind = rs._rs.choice(mn, size=k, replace=False) if k else np.zeros(0, dtype=np.int64)
    j = ind // m
    i = ind - j * m
    if data_rvs is None:
        vals = rs._rs.rand(k)
    else:
        vals = asnumpy(data_rvs(k))
    M = sp.coo_matrix((np.asarray(vals).astype(dt), (i, j)), shape=(m, n)).asformat(format)
    cls = {"csr": _STATE["csr"], "csc": _STATE["csc"], "coo": _STATE["coo"]}[format]
    out = cls(M)
    if format == "csr":
        out._canonical = True  # scipy's coo -> csr sums duplicates and sorts; positions are distinct anyway
    return out


# -------------------------------------------------------------------------- install ----

def _module(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    m.__spmm_b200_compat__ = True
    sys.modules[name] = m
    return m


def install():
    """Register the stand-in modules (idempotent).  A real `cupy` already imported is left alone."""
    if _STATE["installed"]:
        return
    if "cupy" in sys.modules and not getattr(sys.modules["cupy"], "__spmm_b200_compat__", False):
        raise RuntimeError("a real cupy is already imported; spmm_b200.compat must be installed first")
    from . import cusparse as _cs
    nd = _STATE["ndarray"] = _make_ndarray_type()
    csr, csc, coo, rewrap = _wrap_sparse_classes()
    _STATE.update(csr=csr, csc=csc, coo=coo)
    torch = _torch()

    class OutOfMemoryError(MemoryError):  # dense_vs_sparseGEMM/utils.py:158,170
        pass

    memory = _module("cupy.cuda.memory", OutOfMemoryError=OutOfMemoryError, MemoryPool=MemoryPool)
    runtime = _module("cupy.cuda.runtime", memGetInfo=_Runtime.memGetInfo, getDevice=_Runtime.getDevice,
                      getDeviceProperties=_Runtime.getDeviceProperties, deviceSynchronize=_Runtime.deviceSynchronize)
    profiler = _module("cupy.cuda.profiler", start=lambda: None, stop=lambda: None)  # imported, never used (profiler.py:21)
    cuda = _module("cupy.cuda", Stream=Stream, Event=Event, MemoryPool=MemoryPool, PinnedMemoryPool=PinnedMemoryPool,
                   using_allocator=using_allocator, runtime=runtime, memory=memory, profiler=profiler,
                   Device=lambda i=0: types.SimpleNamespace(id=i, synchronize=torch.cuda.synchronize))
    random = _module("cupy.random", seed=_seed, RandomState=RandomState,
                     uniform=lambda low=0.0, high=1.0, size=None, dtype=float: _global_rs().uniform(low, high, size, dtype),
                     standard_normal=lambda size=None, dtype=float: _global_rs().standard_normal(size, dtype),
                     rand=lambda *size: _global_rs().rand(*size))
    default_pool, pinned_pool = MemoryPool.__new__(MemoryPool), PinnedMemoryPool()
    default_pool._base_alloc = default_pool._base_res = default_pool._base_ws = 0
    _module("cupy", cuda=cuda, random=random, ndarray=nd, asarray=asarray, array=asarray, asnumpy=asnumpy,
            abs=lambda x: torch.abs(x).as_subclass(nd), zeros=lambda shape, dtype=np.float32: asarray(np.zeros(shape, _np_dtype(dtype))),
            float32=np.float32, float64=np.float64, float16=np.float16, int32=np.int32, int64=np.int64,
            get_default_memory_pool=lambda: default_pool, get_default_pinned_memory_pool=lambda: pinned_pool,
            __version__="14.0.0a1+spmm_b200.compat")

    def spgemm(a, b, alpha=1, alg=0, chunk_fraction=0.2, verbose=False):  # cupyx/cusparse.py:2007
        return rewrap(_cs.spgemm(a, b, alpha=alpha, alg=alg, chunk_fraction=chunk_fraction, verbose=verbose))

    cusp = _module("cupyx.cusparse", spgemm=spgemm, spmv=lambda *a, **k: rewrap(_cs.spmv(*a, **k)),
                   spmm=lambda *a, **k: rewrap(_cs.spmm(*a, **k)), check_availability=_cs.check_availability)
    sparse = _module("cupyx.scipy.sparse", csr_matrix=csr, csc_matrix=csc, coo_matrix=coo, random=_sparse_random,
                     isspmatrix_csr=lambda x: isinstance(x, csr), issparse=lambda x: isinstance(x, (csr, csc, coo)))
    cscipy = _module("cupyx.scipy", sparse=sparse)
    prof = _module("cupyx.profiler", benchmark=lambda *a, **k: None)  # imported, never called (profiler.py:25)
    _module("cupyx", scipy=cscipy, cusparse=cusp, profiler=prof)
    _STATE["installed"] = True


class _PltStub:
    """matplotlib.pyplot when matplotlib is absent (plots are out of scope):
    every call is a no-op; the cell texts the numerical_error scripts draw are printed at savefig."""

    def __init__(self):
        self.texts = []

    def text(self, x, y, s, *a, **k):
        self.texts.append((x, y, s))

    def savefig(self, path, *a, **k):
        print(f"[matplotlib stub] savefig({path!r}) skipped; cell texts: {self.texts}")

    def __getattr__(self, name):
        return lambda *a, **k: self


def stub_matplotlib():
    try:
        import matplotlib  # noqa: F401
        return False
    except Exception:
        m = types.ModuleType("matplotlib")
        plt = _PltStub()
        pm = types.ModuleType("matplotlib.pyplot")
        pm.__getattr__ = lambda name: getattr(plt, name)
        m.pyplot = pm
        m.use = lambda *a, **k: None
        sys.modules["matplotlib"], sys.modules["matplotlib.pyplot"] = m, pm
        return True


def run(path, argv=()):
    """Run a script file as __main__ with the stand-ins installed (what `python script.py args` does)."""
    install()
    path = os.path.abspath(path)
    old_argv, old_path = sys.argv, list(sys.path)
    sys.argv = [path] + list(argv)
    sys.path.insert(0, os.path.dirname(path))  # dense_vs_sparseGEMM/main.py imports its sibling utils.py
    try:
        runpy.run_path(path, run_name="__main__")
    finally:
        sys.argv, sys.path[:] = old_argv, old_path


def main():
    args = sys.argv[1:]
    if "--stub-matplotlib" in args:
        args.remove("--stub-matplotlib")
        stub_matplotlib()
    if not args:
        print("usage: python -m spmm_b200.compat [--stub-matplotlib] script.py [script args ...]", file=sys.stderr)
        return 2
    run(args[0], args[1:])
    return 0


if __name__ == "__main__":
    os.environ.setdefault("PYTORCH_CUDA_ALLOC_CONF", "expandable_segments:True")
    sys.exit(main())
