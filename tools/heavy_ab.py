#!/usr/bin/env python3
"""Times A @ A on R-MAT inputs (the heavy-row numeric kernel k_num_dense_pieces) and checks the
result bit for bit against the CPU oracle for the scales the oracle finishes in seconds.

  python tools/heavy_ab.py [scale ...] [--f64] [--no-check] [--reps N]

Prints one JSON line per scale: ms per call (median, CUDA events), numeric / symbolic split from
the library's own stats, sha256 of C, and `oracle_equal` (true / false / null when skipped).
Tooling only (it imports oracle/ as the checker, like tests/ do).
"""
import hashlib
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
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    f64 = "--f64" in sys.argv
    check = "--no-check" not in sys.argv
    reps = int(sys.argv[sys.argv.index("--reps") + 1]) if "--reps" in sys.argv else 5
    if "--reps" in sys.argv:
        args = [a for a in args if a != sys.argv[sys.argv.index("--reps") + 1]]
    for sc in [int(a) for a in (args or ["14", "16"])]:
        A = bench.rmat_matrix(sc)
        if f64:
            A = A.astype(np.float64)
        a = spmm_b200.csr_matrix(A)
        a._canonical = True
        c = spmm_b200.cusparse.spgemm(a, a)
        torch.cuda.synchronize()
        ts = []
        for _ in range(reps):
            del c
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            c = spmm_b200.cusparse.spgemm(a, a)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        st = spmm_b200._lib.stats(0)
        Cp, Cj, Cx = c.indptr.cpu().numpy(), c.indices.cpu().numpy(), c.data.cpu().numpy()
        sha = hashlib.sha256(Cp.tobytes() + Cj.tobytes() + Cx.tobytes()).hexdigest()[:16]
        eq = None
        if check and sc <= 16:
            from oracle import oracle
            Op, Oj, Ox = oracle.spgemm((A.indptr, A.indices, A.data), (A.indptr, A.indices, A.data), A.shape[1],
                                       nthreads=oracle.num_threads())
            eq = bool(np.array_equal(Op, Cp) and np.array_equal(Oj, Cj) and
                      np.array_equal(Ox.view(np.uint8), Cx.view(np.uint8)))
        out = {"scale": sc, "dtype": "f64" if f64 else "f32", "ms_median": float(np.median(ts)), "ms_min": float(np.min(ts)), "nnzC": int(Cp[-1]), "sha": sha,
               "oracle_equal": eq}
        if st is not None:
            out["ms_symbolic"] = getattr(st, "ms_symbolic", None)
            out["ms_numeric"] = getattr(st, "ms_numeric", None)
        print(json.dumps(out), flush=True)
        del c, a


if __name__ == "__main__":
    main()
