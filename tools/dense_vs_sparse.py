#!/usr/bin/env python3
"""BASELINE config C4: the reference's dense_vs_sparseGEMM break-even study with libb200sp in the
place of cupyx `A_sparse @ B_sparse` and cuBLAS (torch.matmul, fp32, TF32 off and on) in the place
of cupy `A_dense @ B_dense`.  Same recipe and call shape as the reference
(dense_vs_sparseGEMM/utils.py:238-240 `np.random.seed(seed)` + two independent `scipy.sparse.random`
draws with uniform[0,1) values; :264-275 inputs on the GPU before timing, 1 warm-up + median of
`runs`; main.py:20-66 arguments).  Every sparse result is checked against scipy (indptr/indices
equal, values bitwise) when the product is small enough to do that on the host in seconds.

  python tools/dense_vs_sparse.py --size 8192 --density 1e-5 1e-4 1e-3 1e-2 1e-1 [--runs 5] [--out f.json]
"""
import argparse
import itertools
import json
import os
import sys

import numpy as np
import scipy.sparse as sp_cpu
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import spmm_b200  # noqa: E402
from spmm_b200 import _lib  # noqa: E402


def median_ms(fn, runs, do_warmup=True):
    """utils.py:144-191 (`repeat_gpu`): optional warm-up, then the median of `runs` timed calls;
    timed with CUDA events on the launching stream instead of perf_counter + sync."""
    if do_warmup:
        fn()
    ts = []
    for _ in range(runs):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        out = fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
        del out
    return float(np.sort(ts)[runs // 2])


def run_spmm_case(m, n, p, density, dtype, runs, seed, do_warmup=True, check_limit=3e8):
    np.random.seed(seed)
    A = sp_cpu.random(m, n, density=density, format="csr", dtype=dtype)
    B = sp_cpu.random(n, p, density=density, format="csr", dtype=dtype)
    A.sort_indices()
    B.sort_indices()
    a, b = spmm_b200.csr_matrix(A), spmm_b200.csr_matrix(B)
    products = int(np.diff(B.indptr).astype(np.int64)[A.indices].sum())
    row = {"m": m, "n": n, "p": p, "density": density, "dtype": np.dtype(dtype).name, "nnzA": int(A.nnz), "nnzB": int(B.nnz),
           "products": products, "dense_flops": 2.0 * m * n * p}

    _lib.load().b200sp_reset_stats(_lib.handle(0))
    torch.cuda.reset_peak_memory_stats()
    base = torch.cuda.memory_allocated()
    c = a @ b
    torch.cuda.synchronize()
    row["nnzC"] = int(c.nnz)
    row["sparse_peak_bytes_beyond_inputs"] = int(torch.cuda.max_memory_allocated() - base + _lib.stats(0).peak_workspace_bytes)
    if products <= check_limit:
        R = A @ B
        R.sort_indices()
        ok = (np.array_equal(R.indptr, c.indptr.cpu().numpy()) and np.array_equal(R.indices, c.indices.cpu().numpy())
              and np.array_equal(R.data.view(np.uint32 if dtype == np.float32 else np.uint64),
                                 c.data.cpu().numpy().view(np.uint32 if dtype == np.float32 else np.uint64)))
        row["bitwise_equal_scipy"] = bool(ok)
        assert ok, "SpGEMM result differs from scipy"
    del c
    row["sparse_ms"] = median_ms(lambda: a @ b, runs, do_warmup)
    st = _lib.stats(0)
    row["sparse_ms_symbolic"], row["sparse_ms_numeric"] = float(st.ms_symbolic), float(st.ms_numeric)
    row["sparse_gflops"] = 2.0 * products / row["sparse_ms"] / 1e6

    tdt = torch.float32 if dtype == np.float32 else torch.float64
    Ad = torch.from_numpy(A.toarray()).to("cuda", tdt)
    Bd = torch.from_numpy(B.toarray()).to("cuda", tdt)
    for name, tf32 in (("dense_fp32", False), ("dense_tf32", True)):
        if tdt != torch.float32 and tf32:
            continue
        torch.backends.cuda.matmul.allow_tf32 = tf32
        row[name + "_ms"] = median_ms(lambda: Ad @ Bd, runs, do_warmup)
        row[name + "_tflops"] = row["dense_flops"] / row[name + "_ms"] / 1e9
    torch.backends.cuda.matmul.allow_tf32 = False
    row["dense_peak_bytes_beyond_inputs"] = int(m * p * np.dtype(dtype).itemsize)
    row["speedup_vs_dense_fp32"] = row["dense_fp32_ms"] / row["sparse_ms"]
    del Ad, Bd, a, b
    torch.cuda.empty_cache()

    print("\n" + "*" * 80)
    print("=== b200sp SpGEMM (CSR @ CSR) vs cuBLAS dense GEMM: A @ B ===")
    print(f"A / B shape (CSR) : A=({m}, {n}), B=({n}, {p}), target_density={density}")
    print(f"dtype             : {row['dtype']}   runs: {runs}   products: {products}   nnzC: {row['nnzC']}")
    hdr = f"{'name':40}  {'time(ms)':>10}  {'dPeak Mem (MiB)':>16}"
    print(hdr)
    print("-" * len(hdr))
    print(f"{'A @ B [sparse, inputs_on_gpu]':40}  {row['sparse_ms']:10.4f}  {row['sparse_peak_bytes_beyond_inputs'] / 2**20:16.1f}")
    print(f"{'A @ B [dense fp32, inputs_on_gpu]':40}  {row['dense_fp32_ms']:10.4f}  {row['dense_peak_bytes_beyond_inputs'] / 2**20:16.1f}")
    if "dense_tf32_ms" in row:
        print(f"{'A @ B [dense tf32, inputs_on_gpu]':40}  {row['dense_tf32_ms']:10.4f}  {row['dense_peak_bytes_beyond_inputs'] / 2**20:16.1f}")
    sys.stdout.flush()
    return row


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, nargs="+", required=True)
    ap.add_argument("--density", type=float, nargs="+", required=True)
    ap.add_argument("--dtype", default="float32", choices=["float32", "float64"])
    ap.add_argument("--runs", type=int, default=3)
    ap.add_argument("--seed", type=int, default=42)
    ap.add_argument("--no-warmup", action="store_true")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    dtype = {"float32": np.float32, "float64": np.float64}[args.dtype]
    rows = []
    for size, density in itertools.product(args.size, args.density):
        rows.append(run_spmm_case(size, size, size, density, dtype, args.runs, args.seed, not args.no_warmup))
    if args.out:
        with open(args.out, "w") as f:
            json.dump(rows, f, indent=1)


if __name__ == "__main__":
    main()
