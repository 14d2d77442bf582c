"""Row-panel multi-GPU SpGEMM / SpMV: one process per GPU (torch.distributed), A split into row
panels balanced by intermediate products, B broadcast once over NCCL/NVLink, C row panels landed
in ONE host buffer shared by all ranks (BASELINE.json north_star).

The reference has no multi-GPU path; Gustavson rows are independent, so the only exchange
step is the broadcast of B (one packed buffer: indptr | indices | data - the message shape of
the reference's vendored `_SparseNCCLCommunicator.broadcast`, cupyx/distributed/_nccl_comm.py:651-673,
metadata first, then the arrays).  No reduction.

Flow of `spgemm_row_panels` (every rank):
  1. B (rank 0) is cast to int32 / fp32|fp64, uploaded and broadcast (one `dist.broadcast`);
  2. the rows of A are cut into world * panels_per_rank product-balanced panels dealt round-robin
     (`block_cyclic_panels`), each cut again where its output could exceed the int32 CSR limit;
  3. structure pass of every sub-panel (`cusparse.SpgemmPlan`, one private library handle each);
  4. one all-gather of the per-sub-panel nnz(C): every rank now knows, by an exclusive scan in row
     order, where each of its sub-panels starts in the concatenated C;
  5. value pass per sub-panel into one of two device buffers, D2H on a second stream straight to
     that offset of the shared host buffer (POSIX shared memory mapped and page-locked by every
     rank) while the next sub-panel computes; the rebased indptr slice is written next to it;
  6. barrier; rank 0 holds the whole C as views of the shared buffer.
"""
import atexit
import mmap
import os

import numpy as np

try:  # torch is plumbing only; the partition helpers below work without it
    import torch
    import torch.distributed as dist
except Exception:  # pragma: no cover
    torch = None
    dist = None

INT32_MAX = (1 << 31) - 1


# ------------------------------------------------------------------ partition (host) ----

def row_products(Ap, Aj, Bp):
    """Intermediate products per row of A @ B: sum over the row's nonzeros of nnz(B_k) (int64).
    One multithreaded pass in the library (b200sp_host_row_products; host code, no device needed):
    the numpy form (gather, cumsum, two fancy indexings over nnz(A) int64 values) took 70 - 80 ms for
    the 5.4e6 nonzeros of two C3 panels, more than their C takes to cross PCIe."""
    from . import _lib
    Ap = np.ascontiguousarray(Ap, dtype=np.int32)
    Aj = np.ascontiguousarray(Aj, dtype=np.int32)
    Bp = np.ascontiguousarray(Bp, dtype=np.int32)
    m, k = len(Ap) - 1, len(Bp) - 1
    out = np.empty(max(m, 0), dtype=np.int64)
    st = _lib.load().b200sp_host_row_products(m, k, Ap.ctypes.data, Aj.ctypes.data, Bp.ctypes.data, out.ctypes.data, 0)
    if st != _lib.OK:
        raise ValueError("row_products: malformed indptr or a column index outside B's rows")
    return out


def balanced_row_panels(weights, nparts):
    """Contiguous panels [b[r], b[r+1]) with near-equal total weight.  Greedy with re-targeting:
    panel r ends where the running weight is closest to an equal share of what is LEFT, and takes
    at least one row while rows remain - so a dominant first row (an R-MAT hub) ends its own panel
    and the other parts still share the rest, instead of every boundary collapsing onto it.
    Returns nparts+1 boundaries; panels may be empty when there are fewer rows than parts."""
    w = np.asarray(weights, dtype=np.int64) + 1  # +1: rows without products still cost a visit
    csum = np.cumsum(w)
    nrows = len(w)
    total = int(csum[-1]) if nrows else 0
    bounds, start, base = [0], 0, 0
    for r in range(nparts - 1):
        target = base + (total - base) / (nparts - r)
        i = int(np.searchsorted(csum, target, side="left"))  # csum[i] is the first prefix >= target
        end = i
        if i < nrows:
            below = float(csum[i - 1]) if i > 0 else 0.0
            if float(csum[i]) - target <= target - below:
                end = i + 1
        end = min(max(end, start + 1), nrows)
        bounds.append(end)
        start, base = end, (int(csum[end - 1]) if end > 0 else 0)
    bounds.append(nrows)
    return bounds


def block_cyclic_panels(weights, nparts, per_part=4):
    """nparts * per_part contiguous panels of near-equal weight, dealt round-robin: part r owns
    panels r, r + nparts, ...  Returns one list of (r0, r1) row ranges per part.

    Products per row are the only cost model the host has, and they are not time: on R-MAT the
    hub rows at the top cost twice as much per product as the tail (measured per sub-panel at
    scale 22, tools/panel_cost.py), so one contiguous panel per rank leaves rank 0 1.24x behind
    the mean at 8 GPUs; four interleaved panels per rank bring that to 1.1x."""
    b = balanced_row_panels(weights, nparts * per_part)
    return [[(b[i], b[i + 1]) for i in range(r, nparts * per_part, nparts) if b[i + 1] > b[i]] for r in range(nparts)]


def slice_rows(Ap, Aj, Ax, r0, r1):
    """CSR of rows [r0, r1) with a rebased indptr (views of indices / data)."""
    Ap = np.asarray(Ap)
    s, e = int(Ap[r0]), int(Ap[r1])
    return (Ap[r0:r1 + 1] - Ap[r0]).astype(np.int32), Aj[s:e], Ax[s:e]


def concat_panels(panels):
    """Concatenate C row panels [(indptr, indices, data), ...] with the indptr offset fix-up.
    The result's indptr is int32 when nnz(C) fits, int64 otherwise (a single CSR with more than
    2^31-1 entries cannot have an int32 indptr; the panels themselves always do)."""
    ptrs, off = [np.zeros(1, dtype=np.int64)], 0
    for p, _, _ in panels:
        ptrs.append(np.asarray(p[1:], dtype=np.int64) + off)
        off += int(p[-1])
    ptr = np.concatenate(ptrs)
    if off <= INT32_MAX:
        ptr = ptr.astype(np.int32)
    return (ptr, np.concatenate([j for _, j, _ in panels]), np.concatenate([x for _, _, x in panels]))


def split_rows_by_bound(row_prods, n_cols, r0, r1, limit=INT32_MAX):
    """Cut rows [r0, r1) into consecutive sub-panels whose nnz(C) cannot exceed `limit`:
    nnz(C_i) <= min(products_i, n_cols), so the running sum of that bound decides the cuts.
    Returns the list of boundaries [r0, ..., r1].  A single row always fits (bound <= n_cols)."""
    ub = np.minimum(np.asarray(row_prods[r0:r1], dtype=np.int64), int(n_cols))
    csum = np.cumsum(ub)
    bounds, base = [r0], 0
    while bounds[-1] < r1:
        # last row index whose inclusive running bound stays within base + limit
        k = int(np.searchsorted(csum, base + limit, side="right"))
        k = max(k, bounds[-1] - r0 + 1)  # always take at least one row
        k = min(k, r1 - r0)
        bounds.append(r0 + k)
        base = int(csum[k - 1])
    return bounds


def plan_sub_panels(row_prods, n_cols, world, per_part=4, limit=INT32_MAX):
    """The whole decomposition, identical on every rank: list over ranks of that rank's
    sub-panels [(p0, p1), ...] in ascending row order."""
    out = []
    for ranges in block_cyclic_panels(row_prods, world, per_part):
        subs = []
        for r0, r1 in ranges:
            cuts = split_rows_by_bound(row_prods, n_cols, r0, r1, limit)
            subs.extend(zip(cuts[:-1], cuts[1:]))
        out.append(subs)
    return out


def spgemm_panels(A, b_dev, shape, r0=0, r1=None, limit=INT32_MAX, alg=0, chunk_fraction=0.2, alpha=1.0,
                  row_prods=None):
    """Generator over the C sub-panels of rows [r0, r1) of A @ B on the current GPU.

    A: host CSR triple; b_dev: device csr_matrix of B; yields (p0, p1, c_dev) with c_dev a device
    csr_matrix of rows [p0, p1).  Every sub-panel has fewer than `limit` entries (int32 indptr),
    so a product with more than 2^31-1 nonzeros - R-MAT scale >= 20 squared - streams through one
    GPU panel by panel; the consumer decides what to keep (D2H, checksum, ...)."""
    from . import cusparse
    from .sparse import csr_matrix
    m, k, n = shape
    r1 = m if r1 is None else r1
    Ap, Aj, Ax = A
    if row_prods is None:
        row_prods = row_products(Ap, Aj, b_dev.indptr.cpu().numpy())
    cuts = split_rows_by_bound(row_prods, n, r0, r1, limit)
    for p0, p1 in zip(cuts[:-1], cuts[1:]):
        pp, pj, px = slice_rows(Ap, Aj, Ax, p0, p1)
        a = csr_matrix((np.ascontiguousarray(px), np.ascontiguousarray(pj), pp), shape=(p1 - p0, k))
        a._canonical = True
        yield p0, p1, cusparse.spgemm(a, b_dev, alpha=alpha, alg=alg, chunk_fraction=chunk_fraction)


# ------------------------------------------------------- dynamic panel queue (one node) ---

class PanelQueue:
    """A counter shared by the ranks of one node (8 bytes in /dev/shm, updated under flock): ranks
    draw panel indices as they come free instead of owning a fixed list.  Products are the only cost
    model the host has and they are not time (block_cyclic_panels): with a fixed deal the rank times
    of the C5 job on 8 GPUs spread 68 - 92 ms around a mean of 79.  Collective to create / reset
    (rank 0 writes, everybody passes a barrier); next() is local."""

    def __init__(self):
        import fcntl  # noqa: F401  (POSIX only, like the shared host buffer)
        rank = dist.get_rank()
        name = [None]
        if rank == 0:
            root = "/dev/shm" if os.path.isdir("/dev/shm") else (os.environ.get("TMPDIR") or "/tmp")
            name[0] = os.path.join(root, f"b200sp_q_{os.getpid()}_{np.random.default_rng().integers(1 << 30)}")
            with open(name[0], "wb") as f:
                f.write(b"\0" * 8)
        dist.broadcast_object_list(name, src=0)
        self.path, self.owner = name[0], rank == 0
        self.fd = os.open(self.path, os.O_RDWR)
        dist.barrier()
        if self.owner:
            os.unlink(self.path)  # the descriptors keep it alive

    def reset(self):
        """Collective: the counter is 0 for every rank when this returns."""
        dist.barrier()
        if self.owner:
            os.pwrite(self.fd, (0).to_bytes(8, "little"), 0)
        dist.barrier()

    def next(self):
        import fcntl
        fcntl.flock(self.fd, fcntl.LOCK_EX)
        try:
            v = int.from_bytes(os.pread(self.fd, 8, 0), "little")
            os.pwrite(self.fd, (v + 1).to_bytes(8, "little"), 0)
        finally:
            fcntl.flock(self.fd, fcntl.LOCK_UN)
        return v

    def close(self):
        if self.fd is not None:
            os.close(self.fd)
            self.fd = None


def work_ordered_panels(row_prods, n_cols, nparts, limit=INT32_MAX):
    """nparts contiguous row panels of near-equal products, each cut further so that its output stays
    below `limit` entries, heaviest first (the order in which a PanelQueue hands them out)."""
    b = balanced_row_panels(row_prods, nparts)
    out = []
    for r0, r1 in zip(b[:-1], b[1:]):
        if r1 > r0:
            cuts = split_rows_by_bound(row_prods, n_cols, r0, r1, limit)
            out.extend(zip(cuts[:-1], cuts[1:]))
    csum = np.concatenate(([0], np.cumsum(np.asarray(row_prods, dtype=np.int64))))
    out.sort(key=lambda p: -(int(csum[p[1]]) - int(csum[p[0]])))
    return [(int(a), int(c)) for a, c in out]


# ------------------------------------------------------------------- B broadcast ----------

def cast_csr_host(M, val_dtype=None):
    """(indptr, indices, data) as C-contiguous int32 / int32 / fp32|fp64 host arrays.  SciPy hands
    out int64 index arrays for large or COO-built matrices; the packed broadcast below is sized
    for int32, so everything is cast here and an index that does not fit raises."""
    p, j, x = (M.indptr, M.indices, M.data) if hasattr(M, "indptr") else M
    p, j, x = np.asarray(p), np.asarray(j), np.asarray(x)
    if len(p) and int(p[-1]) > INT32_MAX:
        raise OverflowError("nnz does not fit the int32 indptr of CSR")
    if j.size and j.dtype.itemsize > 4 and (int(j.max()) > INT32_MAX or int(j.min()) < 0):
        raise OverflowError("a column index does not fit int32")
    vdt = np.dtype(val_dtype) if val_dtype is not None else (np.dtype(np.float64) if x.dtype == np.float64 else np.dtype(np.float32))
    if vdt not in (np.float32, np.float64):
        raise ValueError(f"Only float32 and float64 are supported (got {vdt})")
    return (np.ascontiguousarray(p, dtype=np.int32), np.ascontiguousarray(j, dtype=np.int32),
            np.ascontiguousarray(x, dtype=vdt))


def pack_csr(p, j, x):
    """One contiguous byte buffer: indptr | indices | data (single collective, one launch)."""
    if p.dtype != torch.int32 or j.dtype != torch.int32 or x.dtype not in (torch.float32, torch.float64):
        raise TypeError("pack_csr expects int32 indptr / indices and float32 / float64 values")
    return torch.cat([p.contiguous().view(torch.uint8), j.contiguous().view(torch.uint8),
                      x.contiguous().view(torch.uint8)])


def unpack_csr(buf, nrows, nnz, val_dtype):
    o1 = (nrows + 1) * 4
    o2 = o1 + nnz * 4
    return (buf[:o1].view(torch.int32), buf[o1:o2].view(torch.int32), buf[o2:].view(val_dtype))


def broadcast_csr(B, src=0, device=None, meta=None):
    """Broadcast a CSR (tuple of int32 / int32 / fp32|fp64 tensors on `src`, None elsewhere).
    Metadata (rows, nnz, dtype code) goes first as 3 int64, then the packed arrays in one
    dist.broadcast.  `meta` = (rows, nnz, dtype code) when every rank already knows them (an SPMD
    job that multiplies by the same B again): the metadata exchange - a second collective and a
    host synchronisation in front of the payload - is skipped."""
    rank = dist.get_rank()
    device = device or (B[0].device if B is not None else torch.device("cpu"))
    if meta is None:
        mt = torch.zeros(3, dtype=torch.int64, device=device)
        if rank == src:
            p, j, x = B
            mt[0], mt[1], mt[2] = p.numel() - 1, j.numel(), 1 if x.dtype == torch.float64 else 0
        dist.broadcast(mt, src=src)
        meta = mt.tolist()
    nrows, nnz, code = (int(v) for v in meta)
    vdt = torch.float64 if code else torch.float32
    nbytes = (nrows + 1) * 4 + nnz * 4 + nnz * (8 if code else 4)
    if rank == src:
        buf = pack_csr(*B)
        if buf.numel() != nbytes:
            raise ValueError(f"packed CSR has {buf.numel()} bytes, metadata says {nbytes}")
    else:
        buf = torch.empty(nbytes, dtype=torch.uint8, device=device)
    dist.broadcast(buf, src=src)
    return unpack_csr(buf, nrows, nnz, vdt)


# ------------------------------------------------------- the shared host buffer of C ------

class HostArena:
    """One host buffer mapped by every rank of the job: a file in /dev/shm (tmpfs) created by
    rank 0, mmap'ed by all, and - on the CUDA path - page-locked in every process with
    b200sp_host_register, so that each rank's D2H copies land at its offset at DMA speed and the
    concatenation costs nothing.  Grow-only and cached across calls (page-locking gigabytes takes
    longer than the product itself)."""

    def __init__(self):
        self.path, self.mm, self.size, self.base, self.registered, self.owner = None, None, 0, None, False, False
        self.ranges = []      # page-aligned [start, end) byte ranges this process has page-locked
        self.unpinned = False  # a registration failed: copies into the buffer are staged by the driver (slow, correct)

    def _unregister_all(self):
        if self.ranges and self.base is not None:
            from . import _lib
            lib = _lib.load()
            for a, _b in self.ranges:
                lib.b200sp_host_unregister(self.base.ctypes.data + a)
        self.ranges = []

    def register_range(self, offset, nbytes):
        """Page-lock [offset, offset + nbytes) of the buffer for this process (idempotent, cached).  A rank
        registers only what it writes: the whole buffer in every process does not scale - at 8 ranks
        cudaHostRegister of the 17 GB C of eight C3 panels failed in all of them."""
        if not self.registered or nbytes <= 0 or self.unpinned:
            return
        a, b = offset // 4096 * 4096, min(self.size, (offset + nbytes + 4095) // 4096 * 4096)
        for ra, rb in self.ranges:
            if ra <= a and b <= rb:
                return
        # drop cached ranges that overlap the new one (a page cannot be registered twice in one process)
        from . import _lib
        lib = _lib.load()
        keep = []
        for ra, rb in self.ranges:
            if rb <= a or b <= ra:
                keep.append((ra, rb))
            else:
                lib.b200sp_host_unregister(self.base.ctypes.data + ra)
                a, b = min(a, ra), max(b, rb)
        self.ranges = keep
        st = lib.b200sp_host_register(self.base.ctypes.data + a, b - a)
        if st != _lib.OK:
            import warnings
            warnings.warn(f"cudaHostRegister of {b - a} bytes of the shared C buffer failed: copying through pageable memory")
            self.unpinned = True
            return
        self.ranges.append((a, b))

    def _close(self):
        if self.mm is None:
            return
        self._unregister_all()
        self.base = None
        try:
            self.mm.close()
        except BufferError:  # views handed to the caller are still alive; the mapping goes with them
            pass
        if self.owner and self.path and os.path.exists(self.path):
            os.unlink(self.path)
        self.mm, self.size, self.registered = None, 0, False

    def ensure(self, nbytes, rank, register):
        """Collective: all ranks call with the same nbytes."""
        if nbytes <= self.size:
            return
        self._close()
        size = max(4096, (int(nbytes * 1.25) + 4095) & ~4095)
        name = [None]
        if rank == 0:
            root = "/dev/shm" if os.path.isdir("/dev/shm") else (os.environ.get("TMPDIR") or "/tmp")
            name[0] = os.path.join(root, f"b200sp_c_{os.getpid()}_{np.random.default_rng().integers(1 << 30)}")
            fd = os.open(name[0], os.O_CREAT | os.O_RDWR | os.O_EXCL, 0o600)
            os.ftruncate(fd, size)
            os.close(fd)
        dist.broadcast_object_list(name, src=0)
        self.path, self.owner = name[0], rank == 0
        fd = os.open(self.path, os.O_RDWR)
        try:
            self.mm = mmap.mmap(fd, size)
        finally:
            os.close(fd)
        self.size = size
        self.base = np.frombuffer(self.mm, dtype=np.uint8)
        self.registered = bool(register)  # ranges are page-locked on demand (register_range)
        self.unpinned = False
        dist.barrier()  # everyone has the file open: rank 0 may drop the name
        if self.owner:
            os.unlink(self.path)

    def view(self, offset, count, dtype):
        return np.frombuffer(self.mm, dtype=dtype, count=count, offset=offset)


_ARENA = HostArena() if torch is not None else None
if _ARENA is not None:
    atexit.register(_ARENA._close)


def _layout(m, nnz, vsize):
    """Offsets of [indptr int64 (m+1) | indices int32 nnz | data nnz*vsize] in the arena."""
    o_j = ((m + 1) * 8 + 4095) & ~4095
    o_x = (o_j + nnz * 4 + 4095) & ~4095
    return o_j, o_x, o_x + nnz * vsize


# ------------------------------------------------------------------- driver ----------------

class _InjectedEngine:
    """Sub-panel engine around a caller-supplied `compute` (host-logic tests on gloo / CPU)."""

    def __init__(self, compute, Bd, k, n, dev):
        self.compute, self.Bd, self.k, self.n, self.dev, self.res = compute, Bd, k, n, dev, []

    def plan(self, panel):
        t = tuple(torch.from_numpy(np.ascontiguousarray(a)).to(self.dev) for a in panel)
        cp, cj, cx = self.compute(t, self.Bd, (len(panel[0]) - 1, self.k, self.n))
        self.res.append((cp.cpu().numpy(), cj.cpu().numpy(), cx.cpu().numpy()))
        return int(self.res[-1][0][-1])

    def finish(self, i, dst_j, dst_x):
        cp, cj, cx = self.res[i]
        dst_j[:] = cj
        dst_x[:] = cx
        return cp

    def close(self):
        self.res = []


# private library handles of the sub-panel engines, kept across calls per device: creating one costs
# a pinned allocation, a stream and a dozen events, its workspaces a few cudaMalloc / cudaFree pairs
_HANDLE_POOL = {}


def _take_handle(device_index):
    from . import _lib
    pool = _HANDLE_POOL.setdefault(device_index, [])
    return pool.pop() if pool else _lib.new_handle(device_index)


def _give_handle(device_index, h):
    pool = _HANDLE_POOL.setdefault(device_index, [])
    if len(pool) < 16:
        pool.append(h)
    else:
        from . import _lib
        _lib.destroy_handle(h)


class _GpuEngine:
    """Sub-panel engine on libb200sp: structure pass per sub-panel now (private handle, the plan
    stays alive), value pass + D2H later, double-buffered so the copy of sub-panel i runs under
    the kernels of sub-panel i + 1."""

    def __init__(self, Bd, k, n, dev, alg, chunk_fraction, alpha):
        from .sparse import csr_matrix
        self.dev, self.k, self.n, self.alg, self.cf, self.alpha = dev, k, n, alg, chunk_fraction, alpha
        self.b = csr_matrix((Bd[2], Bd[1], Bd[0]), shape=(k, n))
        self.b._canonical = True
        self.plans, self.handles, self.keep = [], [], []
        self.copy_stream = torch.cuda.Stream(device=dev)
        self.bufs, self.buf_free = None, None

    def plan(self, panel):
        from . import _lib, cusparse
        from .sparse import csr_matrix
        pp, pj, px = panel
        a = csr_matrix((np.ascontiguousarray(px), np.ascontiguousarray(pj), pp), shape=(len(pp) - 1, self.k))
        a._canonical = True
        h = _take_handle(self.dev.index)
        self.handles.append(h)
        p = cusparse.SpgemmPlan(a, self.b, alg=self.alg, chunk_fraction=self.cf, handle=h)
        self.plans.append(p)
        return p.nnz

    def finish(self, i, dst_j, dst_x):
        from . import _lib
        p = self.plans[i]
        if self.bufs is None:  # two (indices, data) device buffers of the largest sub-panel
            cap = max(1, max(q.nnz for q in self.plans))
            vdt = self.b.data.dtype
            self.bufs = [(torch.empty(cap, dtype=torch.int32, device=self.dev), torch.empty(cap, dtype=vdt, device=self.dev))
                         for _ in range(2)]
            self.buf_free = [None, None]
        slot = i & 1
        cur = torch.cuda.current_stream(self.dev)
        if self.buf_free[slot] is not None:
            cur.wait_event(self.buf_free[slot])  # the D2H of sub-panel i - 2 has left this buffer
        c = p.numeric(self.alpha, out=self.bufs[slot])
        done = torch.cuda.Event()
        done.record(cur)
        self.copy_stream.wait_event(done)
        lib, h = _lib.load(), self.handles[i]
        sp = self.copy_stream.cuda_stream
        if p.nnz:
            _lib.check(h, lib.b200sp_memcpy_async(h, dst_j.ctypes.data, c.indices.data_ptr(), p.nnz * 4, 1, sp))
            _lib.check(h, lib.b200sp_memcpy_async(h, dst_x.ctypes.data, c.data.data_ptr(), p.nnz * dst_x.itemsize, 1, sp))
        ev = torch.cuda.Event()
        ev.record(self.copy_stream)
        self.buf_free[slot] = ev
        cp = p.indptr.cpu().numpy()  # small; also orders the host behind this sub-panel's kernels
        self.plans[i] = None
        return cp

    def close(self):
        from . import _lib
        self.copy_stream.synchronize()
        torch.cuda.current_stream(self.dev).synchronize()
        for h in self.handles:
            _give_handle(self.dev.index, h)
        self.handles, self.plans, self.bufs = [], [], None


def spgemm_row_panels(A, B, shape, alg=0, chunk_fraction=0.2, alpha=1.0, compute=None, gather=True,
                      panels_per_rank=4, limit=INT32_MAX, copy=False, stats=None):
    """C = alpha * A @ B over all ranks of the default process group.

    A: host CSR triple (numpy), present on every rank (each rank uploads only its panels);
    B: host CSR (scipy matrix or triple) on rank 0 (None elsewhere) - uploaded once and broadcast.
    compute(panel_triple_on_device, B_triple_on_device, (mp, k, n)) -> C triple: tests of the host
    logic may inject one; the default is the libb200sp GPU path.
    gather=True: returns (indptr, indices, data) of the whole C on rank 0 (None elsewhere) - views
    of the shared host buffer, valid until the next call (copy=True: private copies); indptr is
    int32 when nnz(C) fits, int64 otherwise.
    gather=False: returns this rank's [(p0, p1, (indptr, indices, data)), ...] sub-panels (views).
    stats: optional dict that receives sub-panel counts / nnz.
    """
    rank, world = dist.get_rank(), dist.get_world_size()
    m, k, n = (int(v) for v in shape)
    backend = dist.get_backend()
    on_gpu = backend == "nccl"
    dev = torch.device("cuda", torch.cuda.current_device()) if on_gpu else torch.device("cpu")
    Ap, Aj, Ax = A
    Bd = None
    if rank == 0:
        Bd = tuple(torch.from_numpy(a).to(dev) for a in cast_csr_host(B))
    Bd = broadcast_csr(Bd, src=0, device=dev)
    vnp = np.dtype(np.float64) if Bd[2].dtype == torch.float64 else np.dtype(np.float32)
    Bp_host = Bd[0].cpu().numpy()
    row_prods = row_products(Ap, Aj, Bp_host)
    subs_all = plan_sub_panels(row_prods, n, world, panels_per_rank, limit)
    mine = subs_all[rank]
    if compute is None:
        if not on_gpu:
            raise RuntimeError("spgemm_row_panels needs the NCCL backend (one GPU per rank); there is no CPU fallback")
        eng = _GpuEngine(Bd, k, n, dev, alg, chunk_fraction, alpha)
    else:
        eng = _InjectedEngine(compute, Bd, k, n, dev)
    try:
        # ---- structure pass of every sub-panel, then one exchange of the counts ------------
        Ajc, Axc = np.asarray(Aj), np.asarray(Ax, dtype=vnp)  # one cast, the sub-panels are views
        if Ajc.dtype != np.int32:
            Ajc = Ajc.astype(np.int32)
        my_nnz = [eng.plan(slice_rows(Ap, Ajc, Axc, p0, p1)) for p0, p1 in mine]
        width = max(1, max(len(s) for s in subs_all))
        mine_t = torch.zeros(width, dtype=torch.int64, device=dev)
        if my_nnz:
            mine_t[:len(my_nnz)] = torch.tensor(my_nnz, dtype=torch.int64)
        allc = [torch.zeros(width, dtype=torch.int64, device=dev) for _ in range(world)]
        dist.all_gather(allc, mine_t)
        counts = [t.cpu().numpy() for t in allc]
        # exclusive scan over all sub-panels in row order
        order = sorted((p0, r, i) for r, subs in enumerate(subs_all) for i, (p0, _p1) in enumerate(subs))
        offsets, run = {}, 0
        for _p0, r, i in order:
            offsets[(r, i)] = run
            run += int(counts[r][i])
        total = run
        o_j, o_x, nbytes = _layout(m, total, vnp.itemsize)
        _ARENA.ensure(nbytes, rank, register=on_gpu and compute is None)
        ptr = _ARENA.view(0, m + 1, np.int64)
        idx = _ARENA.view(o_j, total, np.int32)
        val = _ARENA.view(o_x, total, vnp)
        if rank == 0:
            ptr[0] = 0
        # ---- value pass + D2H to the sub-panel's offset ------------------------------------
        for i, (p0, p1) in enumerate(mine):  # page-lock what this rank is about to write (cached across calls)
            off, cnt = offsets[(rank, i)], int(my_nnz[i])
            _ARENA.register_range(o_j + off * 4, cnt * 4)
            _ARENA.register_range(o_x + off * vnp.itemsize, cnt * vnp.itemsize)
        for i, (p0, p1) in enumerate(mine):
            off, cnt = offsets[(rank, i)], int(my_nnz[i])
            cp = eng.finish(i, idx[off:off + cnt], val[off:off + cnt])
            ptr[p0 + 1:p1 + 1] = cp[1:].astype(np.int64) + off
    finally:
        eng.close()
    if stats is not None:
        stats.update({"sub_panels": [len(s) for s in subs_all], "nnz": total, "my_nnz": int(sum(my_nnz))})
    dist.barrier()  # all panels are in the buffer
    if not gather:
        out = []
        for i, (p0, p1) in enumerate(mine):
            off, cnt = offsets[(rank, i)], int(my_nnz[i])
            lp = (ptr[p0:p1 + 1] - off).astype(np.int32)
            lp[0] = 0
            out.append((p0, p1, (lp, idx[off:off + cnt], val[off:off + cnt])))
        return out
    if rank != 0:
        return None
    res = (ptr.astype(np.int32) if total <= INT32_MAX else ptr, idx, val)
    return tuple(np.array(a) for a in res) if copy else res


def spmv_row_panels(A, x, shape, compute=None, gather=True):
    """y = A @ x over all ranks: contiguous row panels of A balanced by nonzeros, x broadcast
    from rank 0 (one dist.broadcast), no reduction; y panels all-gathered (padded to the longest
    panel: one collective on the device) and concatenated.

    A: host CSR triple on every rank; x: host vector on rank 0 (None elsewhere).
    compute(panel_triple_on_device, x_on_device, (mp, n)) -> device y panel; default = the GPU path.
    Returns y (numpy, length m) on every rank when gather=True, else (r0, r1, y panel).
    """
    rank, world = dist.get_rank(), dist.get_world_size()
    m, n = shape
    Ap, Aj, Ax = A
    backend = dist.get_backend()
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    vdt = torch.float64 if np.asarray(Ax).dtype == np.float64 else torch.float32
    xd = torch.from_numpy(np.ascontiguousarray(x)).to(device=dev, dtype=vdt).reshape(-1) if rank == 0 else \
        torch.empty(n, dtype=vdt, device=dev)
    if xd.numel() != n:
        raise ValueError("dimension mismatch")
    dist.broadcast(xd, src=0)
    bounds = balanced_row_panels(np.diff(np.asarray(Ap)).astype(np.int64), world)
    r0, r1 = bounds[rank], bounds[rank + 1]
    pp, pj, px = slice_rows(Ap, Aj, Ax, r0, r1)
    panel = tuple(torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (pp, pj, px))
    if compute is None:
        compute = _gpu_spmv
    y = compute(panel, xd, (r1 - r0, n))
    if not gather:
        return r0, r1, y.cpu().numpy()
    width = max(1, max(bounds[r + 1] - bounds[r] for r in range(world)))
    mine = torch.zeros(width, dtype=vdt, device=dev)
    mine[:r1 - r0] = y.to(vdt).reshape(-1)
    parts = [torch.empty(width, dtype=vdt, device=dev) for _ in range(world)]
    dist.all_gather(parts, mine)
    return torch.cat([parts[r][:bounds[r + 1] - bounds[r]] for r in range(world)]).cpu().numpy()


def _gpu_spmv(panel, xd, shape):
    from . import cusparse
    from .sparse import csr_matrix
    mp, n = shape
    a = csr_matrix((panel[2], panel[1], panel[0]), shape=(mp, n))
    a._canonical = True
    return cusparse.spmv(a, xd)
