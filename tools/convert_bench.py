#!/usr/bin/env python3
"""Times the device format conversions on the C2 recipe
(others/spmm/test.py:6-17: randint coordinates, COO -> CSR with duplicates summed) and on the
C3 matrix, with torch's own COO coalesce / to_sparse_csr (a library path on the same box) and
scipy (the reference's CPU path) beside them.  Prints one JSON object.

    python tools/convert_bench.py [--log2n 20] [--density 1e-5] [--reps 20]
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import scipy.sparse as sp
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import spmm_b200  # noqa: E402


def gpu_ms(fn, reps, flush):
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        flush.fill_(1.0)   # evict L2, and keep the GPU busy so that the events bracket GPU time only
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--log2n", type=int, default=20)
    ap.add_argument("--density", type=float, default=1e-5)
    ap.add_argument("--reps", type=int, default=20)
    args = ap.parse_args()
    n = 1 << args.log2n
    k = int(args.density * n * n)
    rng = np.random.default_rng(0)
    row = rng.integers(0, n, k, dtype=np.int32)
    col = rng.integers(0, n, k, dtype=np.int32)
    val = rng.random(k, dtype=np.float32)
    flush = torch.empty(64 << 20, dtype=torch.float32, device="cuda")
    out = {"n": n, "entries": k}

    coo = spmm_b200.coo_matrix((val, (row, col)), shape=(n, n))
    csr = coo.tocsr()
    out["nnz_csr"] = csr.nnz
    # bytes a conversion must move at least once: read 12 B/entry, write 8 B/entry + pointers
    floor = 12 * k + 8 * csr.nnz + 4 * (n + 1)
    ms = gpu_ms(coo.tocsr, args.reps, flush)
    out["coo_to_csr"] = {"ms": ms, "Mentries_per_s": k / ms / 1e3, "floor_bytes": floor, "GBps_of_floor": floor / ms / 1e6}
    csc = csr.tocsc()
    ms = gpu_ms(csc.tocsr, args.reps, flush)
    out["csc_to_csr"] = {"ms": ms, "Mentries_per_s": csr.nnz / ms / 1e3}
    ms = gpu_ms(csr.tocsc, args.reps, flush)
    out["csr_to_csc"] = {"ms": ms}
    other = spmm_b200.csr_matrix((csr.data * 1.0001, csr.indices, csr.indptr), shape=csr.shape)
    other._canonical = True
    ms = gpu_ms(lambda: spmm_b200.max_abs_diff(csr, other), args.reps, flush)
    out["max_abs_diff"] = {"ms": ms, "GBps": 2 * csr.nnz * 8 / ms / 1e6}

    # library path on the same GPU: torch COO coalesce (+ CSR conversion)
    ti = torch.stack([torch.from_numpy(row.astype(np.int64)), torch.from_numpy(col.astype(np.int64))]).cuda()
    tv = torch.from_numpy(val).cuda()

    def torch_path():
        return torch.sparse_coo_tensor(ti, tv, (n, n)).coalesce().to_sparse_csr()
    try:
        out["torch_coalesce_to_csr"] = {"ms": gpu_ms(torch_path, args.reps, flush)}
    except Exception as e:  # noqa: BLE001
        out["torch_coalesce_to_csr"] = {"error": str(e)[:200]}

    # the reference's CPU path
    t0 = time.perf_counter()
    M = sp.coo_matrix((val, (row, col)), shape=(n, n)).tocsr()
    M.sum_duplicates()
    M.sort_indices()
    out["scipy_coo_to_csr_ms"] = (time.perf_counter() - t0) * 1e3
    t0 = time.perf_counter()
    M.tocsc()
    out["scipy_csr_to_csc_ms"] = (time.perf_counter() - t0) * 1e3
    print(json.dumps(out))


if __name__ == "__main__":
    main()
