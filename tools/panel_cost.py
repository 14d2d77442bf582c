#!/usr/bin/env python3
"""Per-sub-panel cost of the streamed R-MAT product on one GPU: rows, sum of products, sum of the
output bound min(products, n), nnz(A), nnz(C) and milliseconds for each sub-panel - the data behind
the row-panel cost model of multi.balanced_row_panels (products alone leave the ranks 24 % apart at
N=8: hub rows and tail rows differ in cost per product).

  python tools/panel_cost.py [scale]  ->  one JSON line per sub-panel, then a least-squares fit
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import spmm_b200  # noqa: E402
from spmm_b200 import multi  # noqa: E402


def main():
    scale = int(sys.argv[1]) if len(sys.argv) > 1 else 22
    A = bench.rmat_matrix(scale)
    m, n = A.shape
    rp = multi.row_products(A.indptr, A.indices, A.indptr)
    b = spmm_b200.csr_matrix(A)
    b._canonical = True
    At = (A.indptr, A.indices, A.data)
    rows = []
    for rep in range(2):  # the first pass warms the allocator and the plans
        rows = []
        gen = multi.spgemm_panels(At, b, (m, n, n), 0, m, row_prods=rp)
        while True:
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            try:
                p0, p1, c = next(gen)
            except StopIteration:
                break
            e1.record()
            torch.cuda.synchronize()
            rows.append({"p0": int(p0), "p1": int(p1), "products": int(rp[p0:p1].sum()),
                         "ub": int(np.minimum(rp[p0:p1], n).sum()), "nnzA": int(A.indptr[p1] - A.indptr[p0]),
                         "nnzC": int(c.nnz), "ms": e0.elapsed_time(e1)})
            del c
    for r in rows:
        print(json.dumps(r))
    X = np.array([[r["products"], r["ub"], r["nnzA"], r["p1"] - r["p0"]] for r in rows], dtype=np.float64)
    y = np.array([r["ms"] for r in rows])
    for cols, name in (([0], "products"), ([0, 1], "products+ub"), ([0, 1, 2], "products+ub+nnzA"), ([0, 1, 2, 3], "+rows")):
        coef, res, *_ = np.linalg.lstsq(X[:, cols], y, rcond=None)
        pred = X[:, cols] @ coef
        print(json.dumps({"fit": name, "coef_ms_per_unit": coef.tolist(),
                          "rel_rms_err": float(np.sqrt(np.mean(((pred - y) / y) ** 2))), "total_ms": float(y.sum())}))


if __name__ == "__main__":
    main()
