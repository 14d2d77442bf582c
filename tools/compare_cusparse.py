#!/usr/bin/env python3
"""Side-by-side: libb200sp vs cuSPARSE ALG1/ALG2/ALG3 on the same device-resident inputs:
time per SpGEMM and peak device memory beyond the inputs (workspace + output), the two axes of
the reference's SpGEMM_alg_comparison (profiler.py:210-230).  Needs a GPU.

  python tools/compare_cusparse.py [--cases c3_1e-4 c3_1e-3 c3_1e-2 c1 c4_1e-2 ...] [--out file.json]
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import spmm_b200  # noqa: E402
from spmm_b200 import _lib  # noqa: E402

EXE = os.path.join(ROOT, "tools", "_build", "cusparse_bench")


def make(case):
    if case == "c1":
        return bench.profiler_matrix(4096, 4096, 1e-3)
    if case.startswith("rmat"):
        return bench.rmat_matrix(int(case[4:]))
    kind, d = case.split("_")
    n = {"c3": 16384, "c4": 8192}[kind]
    return bench.profiler_matrix(n, n, float(d))


def ours(A, alg, reps=7):
    a = spmm_b200.csr_matrix(A)
    a._canonical = True
    ts = []
    _lib.load().b200sp_reset_stats(_lib.handle(0))
    for i in range(reps + 2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        c = spmm_b200.cusparse.spgemm(a, a, alg=alg)
        e1.record()
        torch.cuda.synchronize()
        if i >= 2:
            ts.append(e0.elapsed_time(e1))
    s = _lib.stats(0)
    out_bytes = 4 * (A.shape[0] + 1) + c.nnz * 8
    return {"impl": "b200sp", "alg": alg, "ms_median": float(np.median(ts)), "ms_min": float(min(ts)), "nnzC": c.nnz,
            "peak_bytes_excl_inputs": int(s.peak_workspace_bytes + out_bytes), "workspace_bytes": int(s.peak_workspace_bytes),
            "products": int(s.num_products), "ms_symbolic": s.ms_symbolic, "ms_numeric": s.ms_numeric}


def spmv_case(case, reps):
    """spmv_c2 (BASELINE config C2) or spmv_rmat<scale>: cuSPARSE SpMV (default + CSR_ALG2) against
    our merge-path and vector kernels, L2 flushed before every timed run on both sides."""
    A = bench.spmv_matrix() if case == "spmv_c2" else bench.rmat_matrix(int(case[9:]))
    rows = []
    with tempfile.NamedTemporaryFile(suffix=".bin", delete=False) as f:
        np.array([A.shape[0], A.shape[1], A.nnz], dtype=np.int64).tofile(f)
        A.indptr.astype(np.int32).tofile(f)
        A.indices.astype(np.int32).tofile(f)
        A.data.astype(np.float32).tofile(f)
        path = f.name
    try:
        out = subprocess.run([EXE, path, "0", "0.2", str(max(reps, 10))], capture_output=True, text=True, timeout=600)
        rows.append(json.loads(out.stdout.strip().splitlines()[-1]))
    except Exception as e:
        rows.append({"impl": "cusparse", "op": "spmv", "error": repr(e)})
    os.unlink(path)
    a = spmm_b200.csr_matrix(A)
    a._canonical = True
    x = torch.randn(A.shape[1], device="cuda")
    y = torch.empty(A.shape[0], device="cuda")
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    nbytes = 4 * (A.shape[0] + 1) + A.nnz * 8 + 4 * A.shape[1] + 4 * A.shape[0]
    ours_row = {"impl": "b200sp", "op": "spmv", "m": A.shape[0], "nnz": int(A.nnz), "algorithmic_bytes": nbytes}
    for kern in ("merge", "vector"):
        os.environ["B200SP_SPMV_KERNEL"] = kern
        ts = []
        for i in range(max(reps, 10) + 2):
            flush.fill_(i & 0xff)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            spmm_b200.cusparse.spmv(a, x, y=y)
            e1.record()
            torch.cuda.synchronize()
            if i >= 2:
                ts.append(e0.elapsed_time(e1))
        ours_row[f"ms_{kern}"] = float(np.median(ts))
    os.environ.pop("B200SP_SPMV_KERNEL", None)
    rows.append(ours_row)
    print(f"== {case}: m={A.shape[0]} nnz={A.nnz} algorithmic bytes={nbytes}")
    for r in rows:
        print("  ", {k: v for k, v in r.items()})
    sys.stdout.flush()
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", nargs="+", default=["c3_1e-4", "c3_1e-3", "c3_1e-2"])
    ap.add_argument("--out", default=None)
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    results = {}
    for case in args.cases:
        if case.startswith("spmv"):
            results[case] = spmv_case(case, args.reps)
            continue
        A = make(case)
        rows = []
        with tempfile.NamedTemporaryFile(suffix=".bin", delete=False) as f:
            np.array([A.shape[0], A.shape[1], A.nnz], dtype=np.int64).tofile(f)
            A.indptr.astype(np.int32).tofile(f)
            A.indices.astype(np.int32).tofile(f)
            A.data.astype(np.float32).tofile(f)
            path = f.name
        for alg in (1, 2, 3):
            try:
                out = subprocess.run([EXE, path, str(alg), "0.2", str(args.reps)], capture_output=True, text=True, timeout=600)
                row = json.loads(out.stdout.strip().splitlines()[-1])
                row.setdefault("impl", "cusparse")
                row.setdefault("alg", alg)
                rows.append(row)
            except Exception as e:
                tail = (out.stderr or out.stdout).strip().splitlines()[-1:] if "out" in locals() else []
                rows.append({"impl": "cusparse", "alg": alg, "error": (tail[0] if tail else repr(e))[:200]})
        os.unlink(path)
        for alg in (1, 3):
            rows.append(ours(A, alg))
        results[case] = rows
        prods = rows[-1]["products"]
        print(f"== {case}: n={A.shape[0]} nnzA={A.nnz} products={prods}")
        for r in rows:
            if "error" in r:
                print(f"   {r['impl']:9s} alg{r['alg']}  ERROR {r['error']}")
            else:
                print(f"   {r['impl']:9s} alg{r['alg']}  {r['ms_median']:10.3f} ms  {2 * prods / r['ms_median'] / 1e6:9.1f} GFLOP/s  "
                      f"peak beyond inputs {r['peak_bytes_excl_inputs'] / 2**20:10.1f} MiB  nnzC={r['nnzC']}")
        sys.stdout.flush()
    if args.out:
        with open(args.out, "w") as f:
            json.dump(results, f, indent=1)


if __name__ == "__main__":
    main()
