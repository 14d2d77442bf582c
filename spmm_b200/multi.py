"""Row-panel multi-GPU SpGEMM: one process per GPU (torch.distributed), A split into
contiguous row panels balanced by intermediate products, B broadcast once over NCCL/NVLink,
C row panels concatenated on the host (BASELINE.json north_star).

The reference has no multi-GPU path; Gustavson rows are independent, so the only exchange
step is the broadcast of B (one packed buffer: indptr | indices | data).  No reduction.
"""
import numpy as np

try:  # torch is plumbing only; the partition helpers below work without it
    import torch
    import torch.distributed as dist
except Exception:  # pragma: no cover
    torch = None
    dist = None


# ------------------------------------------------------------------ partition (host) ----

def row_products(Ap, Aj, Bp):
    """Intermediate products per row of A @ B: sum over the row's nonzeros of nnz(B_k)."""
    Ap = np.asarray(Ap, dtype=np.int64)
    blen = np.diff(np.asarray(Bp, dtype=np.int64))
    per_nz = blen[np.asarray(Aj)] if len(Aj) else np.zeros(0, dtype=np.int64)
    csum = np.concatenate(([0], np.cumsum(per_nz, dtype=np.int64)))
    return csum[Ap[1:]] - csum[Ap[:-1]]


def balanced_row_panels(weights, nparts):
    """Contiguous panels [b[r], b[r+1]) with near-equal total weight (prefix-sum split).
    Returns nparts+1 boundaries; panels may be empty when there are fewer rows than parts."""
    w = np.asarray(weights, dtype=np.int64) + 1  # +1: rows without products still cost a visit
    csum = np.cumsum(w)
    total = int(csum[-1]) if len(csum) else 0
    bounds = [0]
    for r in range(1, nparts):
        target = total * r / nparts
        bounds.append(int(np.searchsorted(csum, target, side="left")))
    bounds.append(len(w))
    for r in range(1, len(bounds)):
        bounds[r] = max(bounds[r], bounds[r - 1])
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
    if off <= np.iinfo(np.int32).max:
        ptr = ptr.astype(np.int32)
    return (ptr, np.concatenate([j for _, j, _ in panels]), np.concatenate([x for _, _, x in panels]))


def split_rows_by_bound(row_prods, n_cols, r0, r1, limit=(1 << 31) - 1):
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


def spgemm_panels(A, b_dev, shape, r0=0, r1=None, limit=(1 << 31) - 1, alg=0, chunk_fraction=0.2, alpha=1.0,
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


# ------------------------------------------------------------------- B broadcast ----------

def pack_csr(p, j, x):
    """One contiguous byte buffer: indptr | indices | data (single collective, one launch)."""
    return torch.cat([p.contiguous().view(torch.uint8), j.contiguous().view(torch.uint8),
                      x.contiguous().view(torch.uint8)])


def unpack_csr(buf, nrows, nnz, val_dtype):
    o1 = (nrows + 1) * 4
    o2 = o1 + nnz * 4
    return (buf[:o1].view(torch.int32), buf[o1:o2].view(torch.int32), buf[o2:].view(val_dtype))


def broadcast_csr(B, src=0, device=None):
    """Broadcast a CSR (tuple of tensors on `src`, None elsewhere).  Metadata (rows, nnz, dtype
    code) goes first as 3 int64, then the packed arrays in one dist.broadcast."""
    rank = dist.get_rank()
    device = device or (B[0].device if B is not None else torch.device("cpu"))
    meta = torch.zeros(3, dtype=torch.int64, device=device)
    if rank == src:
        p, j, x = B
        meta[0], meta[1], meta[2] = p.numel() - 1, j.numel(), 1 if x.dtype == torch.float64 else 0
    dist.broadcast(meta, src=src)
    nrows, nnz, code = (int(v) for v in meta.tolist())
    vdt = torch.float64 if code else torch.float32
    nbytes = (nrows + 1) * 4 + nnz * 4 + nnz * (8 if code else 4)
    buf = pack_csr(*B) if rank == src else torch.empty(nbytes, dtype=torch.uint8, device=device)
    dist.broadcast(buf, src=src)
    return unpack_csr(buf, nrows, nnz, vdt)


# ------------------------------------------------------------------- driver ----------------

def spgemm_row_panels(A, B, shape, alg=0, chunk_fraction=0.2, alpha=1.0, compute=None, gather=True):
    """C = alpha * A @ B over all ranks of the default process group.

    A: host CSR triple (numpy), present on every rank (each rank uploads only its panel);
    B: host CSR triple on rank 0 (None elsewhere) - it is uploaded once and broadcast.
    compute(panel_triple_on_device, B_triple_on_device, (mp, k, n)) -> device C triple; the
    default is the libb200sp GPU path (tests of the host logic may inject another).
    Returns the concatenated host C on rank 0 when gather=True (None elsewhere), else this
    rank's (r0, r1, host C panel).
    """
    rank, world = dist.get_rank(), dist.get_world_size()
    m, k, n = shape
    Ap, Aj, Ax = A
    backend = dist.get_backend()
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    Bd = None
    if rank == 0:
        Bd = tuple(torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in B)
    Bd = broadcast_csr(Bd, src=0, device=dev)
    Bp_host = Bd[0].cpu().numpy()
    bounds = balanced_row_panels(row_products(Ap, Aj, Bp_host), world)
    r0, r1 = bounds[rank], bounds[rank + 1]
    pp, pj, px = slice_rows(Ap, Aj, Ax, r0, r1)
    panel = tuple(torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (pp, pj, px))
    if compute is None:
        compute = _gpu_compute(alg, chunk_fraction, alpha)
    cp, cj, cx = compute(panel, Bd, (r1 - r0, k, n))
    host_c = (cp.cpu().numpy(), cj.cpu().numpy(), cx.cpu().numpy())
    if not gather:
        return r0, r1, host_c
    out = [None] * world if rank == 0 else None
    dist.gather_object(host_c, out, dst=0)
    return concat_panels(out) if rank == 0 else None


def _gpu_compute(alg, chunk_fraction, alpha):
    from . import cusparse
    from .sparse import csr_matrix

    def run(panel, Bd, shape):
        mp, k, n = shape
        a = csr_matrix((panel[2], panel[1], panel[0]), shape=(mp, k))
        b = csr_matrix((Bd[2], Bd[1], Bd[0]), shape=(k, n))
        a._canonical = b._canonical = True
        c = cusparse.spgemm(a, b, alpha=alpha, alg=alg, chunk_fraction=chunk_fraction)
        return c.indptr, c.indices, c.data
    return run


def spmv_row_panels(A, x, shape, compute=None, gather=True):
    """y = A @ x over all ranks: contiguous row panels of A balanced by nonzeros, x broadcast
    from rank 0 (one dist.broadcast), no reduction; y panels concatenated on rank 0
    (SURVEY.md section 8e).

    A: host CSR triple on every rank; x: host vector on rank 0 (None elsewhere).
    compute(panel_triple_on_device, x_on_device, (mp, n)) -> device y panel; default = the GPU path.
    Returns y (numpy, length m) on rank 0 when gather=True (None elsewhere), else (r0, r1, y panel).
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
    y = compute(panel, xd, (r1 - r0, n)).cpu().numpy()
    if not gather:
        return r0, r1, y
    out = [None] * world if rank == 0 else None
    dist.gather_object(y, out, dst=0)
    return np.concatenate(out) if rank == 0 else None


def _gpu_spmv(panel, xd, shape):
    from . import cusparse
    from .sparse import csr_matrix
    mp, n = shape
    a = csr_matrix((panel[2], panel[1], panel[0]), shape=(mp, n))
    a._canonical = True
    return cusparse.spmv(a, xd)
