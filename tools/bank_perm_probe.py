#!/usr/bin/env python3
"""Probe: how much do the shared-memory accumulator kernels gain when every aligned group of 32
entries of a B row hits 32 different banks?  B's rows are permuted on the host (the order inside a
B row does not change any bit of C: its columns are distinct), then the same SpGEMM is timed.

  python tools/bank_perm_probe.py [n] [density]
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import spmm_b200  # noqa: E402
from spmm_b200 import _lib  # noqa: E402


def bank_deal(indptr, indices, key_bits=5):
    """dest position of every entry: entries of a row sorted by bank, then dealt like cards over
    the row's aligned 32-blocks (the last block may be short)."""
    m = len(indptr) - 1
    L = np.diff(indptr).astype(np.int64)
    row = np.repeat(np.arange(m), L)
    bank = indices & ((1 << key_bits) - 1)
    order = np.lexsort((indices, bank, row))  # by row, then bank, then column
    start = np.repeat(indptr[:-1].astype(np.int64), L)
    r = np.arange(len(indices)) - start  # rank inside the row of the entry at sorted position
    Lr = np.repeat(L, L)
    G = (Lr + 31) // 32
    rem = Lr - 32 * (G - 1)
    first = r < rem * G
    c = np.where(first, r // G, 0)
    g = np.where(first, r % G, 0)
    r2 = r - rem * G
    Gm = np.maximum(G - 1, 1)
    c = np.where(first, c, rem + r2 // Gm)
    g = np.where(first, g, r2 % Gm)
    dest = start + 32 * g + c
    src = np.empty_like(order)
    src[dest] = order
    return src


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
    d = float(sys.argv[2]) if len(sys.argv) > 2 else 1e-2
    A = bench.profiler_matrix(n, n, d)
    src = bank_deal(A.indptr, A.indices)
    Bp = A.copy()
    Bp.indices = A.indices[src].copy()
    Bp.data = A.data[src].copy()
    # sanity: same rows
    assert (np.sort(Bp.indices[A.indptr[5]:A.indptr[6]]) == A.indices[A.indptr[5]:A.indptr[6]]).all()
    a = spmm_b200.csr_matrix(A)
    a._canonical = True
    b = spmm_b200.csr_matrix(Bp)
    b._canonical = True
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ref = None
    for name, bb in (("sorted B", a), ("bank-dealt B", b), ("sorted B", a), ("bank-dealt B", b)):
        sym, num = [], []
        for i in range(8):
            flush.fill_(i)
            c = spmm_b200.cusparse.spgemm(a, bb)
            st = _lib.stats(0)
            if i >= 3:
                sym.append(st.ms_symbolic)
                num.append(st.ms_numeric)
        print(f"{name:14s} symbolic {np.mean(sym):.4f} ms  numeric {np.mean(num):.4f} ms", flush=True)
        if ref is None:
            ref = c
        else:
            assert torch.equal(ref.indptr, c.indptr) and torch.equal(ref.indices, c.indices)
            assert torch.equal(ref.data.view(torch.int32), c.data.view(torch.int32)), "values differ"
    print("bitwise equal")


if __name__ == "__main__":
    main()
