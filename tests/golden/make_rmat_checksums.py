#!/usr/bin/env python3
"""Golden checksums of the C5 workload (R-MAT A @ A, bench.rmat_matrix) at scales whose output
is too large to commit or even to hold next to the inputs (scale 20: 9.7e9 entries, 78 GB).

Produced on the CPU by the reference's own CPU path (scipy.sparse `A[r0:r1] @ A`, sorted) at
scale 16 and 18 and - scipy and the oracle agreeing bit for bit there - by the multi-threaded
oracle (oracle/spgemm_oracle.c) at scale 20, row chunk by row chunk.  The checksums are
independent of how the rows are partitioned, so a single-GPU run, a streamed sub-panel run and
an N-GPU row-panel run must all reproduce them exactly (arithmetic mod 2^64):

  nnz      number of stored entries of C
  rowsum   sum_i (i + 1) * nnz(C_i)
  colsum   sum_k (col_k + 1)
  colval   sum_k (col_k + 1) * (bits_k + 1),  bits_k = the fp32 value's bit pattern as uint32

Run:  python tests/golden/make_rmat_checksums.py [scales...]     (default 16 18 20)
"""
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(HERE))
import bench  # noqa: E402
from oracle import oracle  # noqa: E402

M64 = (1 << 64) - 1


def chunk_checksums(r0, Cp, Cj, Cx):
    """(nnz, rowsum, colsum, colval) of one row chunk starting at global row r0 (mod 2^64)."""
    with np.errstate(over="ignore"):
        cnt = np.diff(Cp).astype(np.uint64)
        rows = np.arange(r0 + 1, r0 + len(cnt) + 1, dtype=np.uint64)
        col1 = Cj.astype(np.uint64) + np.uint64(1)
        bits1 = Cx.view(np.uint32).astype(np.uint64) + np.uint64(1)
        return (int(Cp[-1]), int((rows * cnt).sum(dtype=np.uint64)), int(col1.sum(dtype=np.uint64)),
                int((col1 * bits1).sum(dtype=np.uint64)))


def combine(parts):
    tot = [0, 0, 0, 0]
    for p in parts:
        for i in range(4):
            tot[i] = (tot[i] + p[i]) & M64
    return {"nnz": tot[0], "rowsum": tot[1], "colsum": tot[2], "colval": tot[3]}


def row_chunks(A, limit=400_000_000):
    from spmm_b200 import multi
    rp = multi.row_products(A.indptr, A.indices, A.indptr)
    cuts = multi.split_rows_by_bound(rp, A.shape[1], 0, A.shape[0], limit)
    return list(zip(cuts[:-1], cuts[1:]))


def via_oracle(A):
    T = (A.indptr, A.indices, A.data)
    parts = []
    for r0, r1 in row_chunks(A):
        P = A[r0:r1]
        Cp, Cj, Cx = oracle.spgemm((P.indptr, P.indices, P.data), T, A.shape[1], nthreads=0)
        parts.append(chunk_checksums(r0, Cp, Cj, Cx))
    return combine(parts)


def via_scipy(A):
    parts = []
    for r0, r1 in row_chunks(A):
        C = (A[r0:r1] @ A).tocsr()
        C.sort_indices()
        parts.append(chunk_checksums(r0, C.indptr, C.indices, C.data))
    return combine(parts)


def main():
    scales = [int(s) for s in sys.argv[1:]] or [16, 18, 20]
    path = os.path.join(HERE, "rmat_checksums.json")
    out = json.load(open(path)) if os.path.exists(path) else {}
    import scipy
    out["_recipe"] = "bench.rmat_matrix(scale): edge factor 16, (0.57,0.19,0.19,0.05), seed 1, fp32 uniform[0,1) values; C = A @ A"
    out["_scipy"] = scipy.__version__
    for sc in scales:
        t = time.time()
        A = bench.rmat_matrix(sc)
        o = via_oracle(A)
        entry = {"n": int(A.shape[0]), "nnzA": int(A.nnz), "products": bench.count_products(A, A), **o, "by": "oracle"}
        if sc <= 18:
            s = via_scipy(A)
            assert s == o, f"scale {sc}: scipy and the oracle disagree: {s} vs {o}"
            entry["by"] = "scipy.sparse (the oracle reproduces it bit for bit)"
        out[f"rmat{sc}"] = entry
        print(sc, entry, f"{time.time() - t:.1f}s", flush=True)
        with open(path, "w") as f:
            json.dump(out, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
