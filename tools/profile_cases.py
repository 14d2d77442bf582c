#!/usr/bin/env python3
"""Runs each headline case a few times so that `ncu` can list / capture the kernels.

  python tools/profile_cases.py [case ...]   cases: c3_1e-4 c3_1e-3 c3_1e-2 c1 spmv c4_1e-1 rmat16
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import spmm_b200  # noqa: E402


def main():
    cases = sys.argv[1:] or ["c3_1e-4", "c3_1e-3", "c3_1e-2", "c1", "spmv"]
    reps = int(os.environ.get("REPS", "2"))
    for c in cases:
        if c == "spmv":
            M = bench.spmv_matrix()
            a = spmm_b200.csr_matrix(M)
            a._canonical = True
            x = torch.randn(M.shape[1], device="cuda")
            for _ in range(reps):
                spmm_b200.cusparse.spmv(a, x)
        else:
            if c.startswith("rmat"):
                A = bench.rmat_matrix(int(c[4:]))
            elif c == "c1":
                A = bench.profiler_matrix(4096, 4096, 1e-3)
            else:
                kind, d = c.split("_")
                n = 16384 if kind == "c3" else 8192
                A = bench.profiler_matrix(n, n, float(d))
            a = spmm_b200.csr_matrix(A)
            a._canonical = True
            for _ in range(reps):
                cc = spmm_b200.cusparse.spgemm(a, a)
            del cc
        torch.cuda.synchronize()
        print("done", c, flush=True)


if __name__ == "__main__":
    main()
