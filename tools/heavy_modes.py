#!/usr/bin/env python3
"""Heavy-row kernel choices on R-MAT A @ A, side by side: the compact rank accumulator
(k_num_rank_cta) at several accumulator sizes, the dense column blocks of round 1
(k_num_dense_pieces), and the mix.  Every run is checked against the CPU-made golden checksums
(tests/golden/rmat_checksums.json).  Scale >= 20 streams sub-panels (multi.spgemm_panels).

  python tools/heavy_modes.py [scale ...] [--reps N] [--configs name,name,...]
One JSON line per (scale, config)."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import spmm_b200  # noqa: E402
from spmm_b200 import _lib, multi  # noqa: E402

CONFIGS = {
    "blocked": {"B200SP_HEAVY": "blocked"},
    "rank96": {"B200SP_HEAVY": "rank", "B200SP_RANK_ACC_KB": "96"},
    "rank64": {"B200SP_HEAVY": "rank", "B200SP_RANK_ACC_KB": "64"},
    "rank48": {"B200SP_HEAVY": "rank", "B200SP_RANK_ACC_KB": "48"},
    "rank192": {"B200SP_HEAVY": "rank", "B200SP_RANK_ACC_KB": "192"},
    "rank96_t512": {"B200SP_HEAVY": "rank", "B200SP_RANK_ACC_KB": "96", "B200SP_HEAVY_MIN": "512"},
    "rank96_t2048": {"B200SP_HEAVY": "rank", "B200SP_RANK_ACC_KB": "96", "B200SP_HEAVY_MIN": "2048"},
    "mixed8": {"B200SP_HEAVY": "mixed", "B200SP_RANK_ACC_KB": "96", "B200SP_HEAVY_DENSE_DIV": "8"},
    "mixed4": {"B200SP_HEAVY": "mixed", "B200SP_RANK_ACC_KB": "96", "B200SP_HEAVY_DENSE_DIV": "4"},
    # short-segment (window) path of the blocked kernel: windows per hand-off turn, products per piece below which it is taken
    "blocked_w1": {"B200SP_HEAVY": "blocked", "B200SP_WINDOW_TURN": "1"},
    "blocked_w4": {"B200SP_HEAVY": "blocked", "B200SP_WINDOW_TURN": "4"},
    "blocked_w8": {"B200SP_HEAVY": "blocked", "B200SP_WINDOW_TURN": "8"},
    "blocked_w4_f64": {"B200SP_HEAVY": "blocked", "B200SP_WINDOW_TURN": "4", "B200SP_FLAT_BELOW": "64"},
    "blocked_w8_f64": {"B200SP_HEAVY": "blocked", "B200SP_WINDOW_TURN": "8", "B200SP_FLAT_BELOW": "64"},
    "blocked_w8_f96": {"B200SP_HEAVY": "blocked", "B200SP_WINDOW_TURN": "8", "B200SP_FLAT_BELOW": "96"},
    "blocked_w8_f128": {"B200SP_HEAVY": "blocked", "B200SP_WINDOW_TURN": "8", "B200SP_FLAT_BELOW": "128"},
    # CTA shape of the blocked kernel: threads per CTA (= longest piece per barrier step) and accumulator kilobytes per CTA
    "nt128_kb100": {"B200SP_HEAVY": "blocked", "B200SP_BLOCKED_NT": "128", "B200SP_BLOCKED_KB": "100"},
    "nt128_kb64": {"B200SP_HEAVY": "blocked", "B200SP_BLOCKED_NT": "128", "B200SP_BLOCKED_KB": "64"},
    "nt128_kb48": {"B200SP_HEAVY": "blocked", "B200SP_BLOCKED_NT": "128", "B200SP_BLOCKED_KB": "48"},
    "nt256_kb48": {"B200SP_HEAVY": "blocked", "B200SP_BLOCKED_NT": "256", "B200SP_BLOCKED_KB": "48"},
    "nt256_kb200": {"B200SP_HEAVY": "blocked", "B200SP_BLOCKED_NT": "256", "B200SP_BLOCKED_KB": "200"},
    # one warp per (row, narrow block): k_num_warp_blocks, accumulator columns per warp
    "warp": {"B200SP_HEAVY": "warp"},
    "warp_ww1280": {"B200SP_HEAVY": "warp", "B200SP_WARP_WW": "1280"},
    "warp_ww5120": {"B200SP_HEAVY": "warp", "B200SP_WARP_WW": "5120"},
    "warp_t512": {"B200SP_HEAVY": "warp", "B200SP_HEAVY_MIN": "512"},
}
KEYS = ("B200SP_WARP_WW", "B200SP_BLOCKED_NT", "B200SP_BLOCKED_KB", "B200SP_HEAVY", "B200SP_RANK_ACC_KB", "B200SP_HEAVY_MIN", "B200SP_HEAVY_DENSE_DIV", "B200SP_WINDOW_TURN", "B200SP_FLAT_BELOW")


def run_once(A, b, rp, dev):
    m, n = A.shape
    chk = torch.zeros(4, dtype=torch.int64, device=dev)
    ms_sym = ms_num = 0.0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_chk = 0.0
    e0.record()
    for p0, _p1, c in multi.spgemm_panels((A.indptr, A.indices, A.data), b, (m, n, n), row_prods=rp):
        st = _lib.stats(dev.index)
        ms_sym += st.ms_symbolic
        ms_num += st.ms_numeric
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record()
        chk += bench.device_checksums(torch, p0, c)
        c1.record()
        torch.cuda.synchronize()
        t_chk += c0.elapsed_time(c1)
        del c
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) - t_chk, ms_sym, ms_num, {k: int(v) & bench.M64 for k, v in zip(("nnz", "rowsum", "colsum", "colval"), chk.tolist())}


def main():
    argv = sys.argv[1:]
    reps = int(argv[argv.index("--reps") + 1]) if "--reps" in argv else 2
    names = argv[argv.index("--configs") + 1].split(",") if "--configs" in argv else list(CONFIGS)
    scales = [int(a) for a in argv if a.isdigit() and (argv.index(a) == 0 or not argv[argv.index(a) - 1].startswith("--"))] or [16, 18]
    dev = torch.device("cuda", 0)
    for sc in scales:
        A = bench.rmat_matrix(sc)
        rp = multi.row_products(A.indptr, A.indices, A.indptr)
        prods = int(rp.sum())
        b = spmm_b200.csr_matrix(A)
        b._canonical = True
        gold = bench.golden_rmat(sc)
        for name in names:
            for k in KEYS:
                os.environ.pop(k, None)
            os.environ.update(CONFIGS[name])
            try:
                run_once(A, b, rp, dev)
                res = [run_once(A, b, rp, dev) for _ in range(reps)]
                ms = float(np.median([r[0] for r in res]))
                got = res[-1][3]
                out = {"scale": sc, "config": name, "ms": ms, "gflops": 2 * prods / ms / 1e6, "ms_symbolic": res[-1][1], "ms_numeric": res[-1][2],
                       "checksums_ok": (all(got[k] == gold[k] for k in got) if gold else None), "build": bench.build_info().get("git")}
            except Exception as e:  # keep going: one bad config must not hide the others
                out = {"scale": sc, "config": name, "error": repr(e)[:300]}
            print(json.dumps(out), flush=True)
        del b
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
