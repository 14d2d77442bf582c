#!/usr/bin/env python3
"""SpMV probe: C2 (uniform random columns) against a banded matrix of the same shape and nnz.
The banded case gathers x from L1/L2-friendly addresses, so the difference between the two is
the cost of the random 4-byte gathers (32-byte L2 sectors), not of streaming A."""
import os
import sys

import numpy as np
import scipy.sparse as sp
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import spmm_b200  # noqa: E402


def timed(fn, reps=20, do_flush=True):
    flush = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")  # long enough for the host to run ahead
    fn(); fn()
    ts = []
    for i in range(reps):
        if do_flush:
            flush.fill_(i)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts)), float(np.min(ts))


def main():
    n = 1 << 20
    M = bench.spmv_matrix()
    per = np.diff(M.indptr)
    rows = np.repeat(np.arange(n, dtype=np.int64), per)
    off = np.concatenate([np.arange(k) for k in per]) if False else (np.arange(M.nnz) - np.repeat(M.indptr[:-1], per))
    cols = (rows + off * 3) % n
    Bd = sp.csr_matrix((M.data, (rows, cols)), shape=(n, n))
    Bd.sum_duplicates(); Bd.sort_indices()
    x = torch.randn(n, device="cuda")
    y = torch.empty(n, device="cuda")
    R = bench.rmat_matrix(20)
    L = bench.profiler_matrix(2000, 1 << 20, 0.005, seed=3)
    for name, A in (("random", M), ("banded", Bd), ("rmat20", R), ("longrow", L)):
        a = spmm_b200.csr_matrix(A); a._canonical = True
        nbytes = 4 * (A.shape[0] + 1) + A.nnz * 8 + 4 * A.shape[1] + 4 * A.shape[0]
        x = torch.randn(A.shape[1], device="cuda")
        y = torch.empty(A.shape[0], device="cuda")
        for kern, pf in (("stream", "1"), ("vector", "1")):
            os.environ["B200SP_SPMV_KERNEL"] = kern
            os.environ["B200SP_SPMV_PREFETCH"] = pf
            med, mn = timed(lambda: spmm_b200.cusparse.spmv(a, x, y=y))
            medw, mnw = timed(lambda: spmm_b200.cusparse.spmv(a, x, y=y), do_flush=False)
            print(f"{name:7s} {kern:7s} nnz={A.nnz} maxrow={int(np.diff(A.indptr).max())} cold median {med*1e3:7.1f} us min {mn*1e3:7.1f} | L2-warm median {medw*1e3:7.1f} min {mnw*1e3:7.1f} us | {nbytes/med/1e6:7.0f} GB/s cold")
    # pure streaming reference: torch copy of the same number of bytes
    src = torch.empty(nbytes // 8, dtype=torch.float32, device="cuda"); dst = torch.empty_like(src)
    med, mn = timed(lambda: dst.copy_(src))
    print(f"torch copy of {src.numel()*4} B (read+write {src.numel()*8} B): median {med*1e3:.1f} us -> {src.numel()*8/med/1e6:.0f} GB/s")


if __name__ == "__main__":
    main()
