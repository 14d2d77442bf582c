#!/usr/bin/env python3
"""SpMV kernels against the row length: uniform random square matrices with 4 .. 256 entries per
row (1.1e7 nonzeros each), every kernel variant, L2 flushed before each run.  Decides the
thresholds of the kernel choice in run_spmv (b200sp_api.cu).  One JSON line per matrix."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import spmm_b200  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    nnz = 11_000_000
    for avg in (4, 8, 12, 16, 24, 32, 64, 128, 256):
        n = nnz // avg
        M = bench.profiler_matrix(n, n, avg / n, seed=avg)
        a = spmm_b200.csr_matrix(M)
        a._canonical = True
        x = torch.randn(n, device=dev)
        y = torch.empty(n, device=dev)
        row = {"avg_row": avg, "n": n, "nnz": int(M.nnz)}
        for kern in ("vector4", "vector+scalar", "merge", "merge+scalar"):
            os.environ["B200SP_SPMV_KERNEL"] = kern
            ts = []
            for i in range(14):
                flush.fill_(i)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                spmm_b200.cusparse.spmv(a, x, y=y)
                e1.record()
                torch.cuda.synchronize()
                if i >= 4:
                    ts.append(e0.elapsed_time(e1))
            row[kern] = float(np.median(ts)) * 1e3
        print(json.dumps(row), flush=True)
        del a


if __name__ == "__main__":
    main()
