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


def slice_rows(Ap, Aj, Ax, r0, r1):
    """CSR of rows [r0, r1) with a rebased indptr (views of indices / data)."""
    Ap = np.asarray(Ap)
    s, e = int(Ap[r0]), int(Ap[r1])
    return (Ap[r0:r1 + 1] - Ap[r0]).astype(np.int32), Aj[s:e], Ax[s:e]


def concat_panels(panels):
    """Concatenate C row panels [(indptr, indices, data), ...] with the indptr offset fix-up."""
    ptrs, off = [np.zeros(1, dtype=np.int64)], 0
    for p, _, _ in panels:
        ptrs.append(np.asarray(p[1:], dtype=np.int64) + off)
        off += int(p[-1])
    if off > np.iinfo(np.int32).max:
        raise OverflowError("nnz(C) does not fit int32 indptr")
    return (np.concatenate(ptrs).astype(np.int32), np.concatenate([j for _, j, _ in panels]),
            np.concatenate([x for _, _, x in panels]))


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
