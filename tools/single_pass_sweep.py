#!/usr/bin/env python3
"""Single-pass against two-pass SpGEMM over the density sweeps of C3 (16384^2) and C4 (8192^2):
where does skipping the structure pass pay?  Decides single_pass_pays() in b200sp_api.cu.
One JSON line per (config, density): products / (m n), both times (median of 7, L2 flushed), bits equal."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import spmm_b200  # noqa: E402
from spmm_b200 import _lib  # noqa: E402


def timed(a, flush, reps=7):
    ts = []
    c = None
    for i in range(reps + 2):
        flush.fill_(i)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        c = spmm_b200.cusparse.spgemm(a, a)
        e1.record()
        torch.cuda.synchronize()
        if i >= 2:
            ts.append(e0.elapsed_time(e1))
    st = _lib.stats(0)
    return float(np.median(ts)), st.ms_symbolic, st.ms_numeric, c


def main():
    dev = torch.device("cuda", 0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    cases = [("c3", 16384, d) for d in (1e-3, 2e-3, 3e-3, 5e-3, 7e-3, 1e-2)] + \
            [("c4", 8192, d) for d in (3e-3, 5e-3, 7e-3, 1e-2, 1.5e-2, 2e-2, 3e-2, 1e-1)]
    for name, n, d in cases:
        A = bench.profiler_matrix(n, n, d)
        a = spmm_b200.csr_matrix(A)
        a._canonical = True
        os.environ["B200SP_SINGLE_PASS"] = "0"
        os.environ.pop("B200SP_SINGLE_PASS_RATIO", None)
        t2, s2, n2, c2 = timed(a, flush)
        os.environ.pop("B200SP_SINGLE_PASS", None)
        os.environ["B200SP_SINGLE_PASS_RATIO"] = "0"
        t1, s1, n1, c1 = timed(a, flush)
        same = bool(torch.equal(c1.indptr, c2.indptr) and torch.equal(c1.indices, c2.indices) and
                    torch.equal(c1.data.view(torch.int32), c2.data.view(torch.int32)))
        prods = bench.count_products(A, A)
        print(json.dumps({"case": name, "n": n, "density": d, "products": int(prods), "products_per_mn": prods / (n * n),
                          "nnzC": int(c1.nnz), "two_pass_ms": t2, "two_pass_sym_ms": s2, "two_pass_num_ms": n2,
                          "single_pass_ms": t1, "single_pass_estimate_ms": s1, "single_pass_kernel_ms": n1,
                          "bits_equal": same}), flush=True)
        del a, c1, c2


if __name__ == "__main__":
    main()
