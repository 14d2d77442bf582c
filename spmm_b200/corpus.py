"""txt-CSR corpus generator: the CuPy-free restatement of the reference's
cupy_cusparse/gen_and_save_alg{1,2,3}_txt.py.

For every (n, density) it draws A and B (seeds s and s+1, rows sorted: gen_and_save_alg1_txt.py:16-24),
computes C with the Python operator `spmm_b200.cusparse.spgemm(A, B, alg=k[, chunk_fraction])`
(:26-27; alg3 variant gen_and_save_alg3_txt.py:27-28) and writes the three matrices as
`<outdir>/{A,B,C_py}_n{N}_dens{d with . -> p}_alg{k}[_cf…]_{indptr,indices,data}.txt`
("%d" / "%.9g", :8-14, :29-33) with the library's writer.  The random draw uses scipy's
generator on the host (CuPy's device RNG is not available); the product runs on the GPU.
"""
import argparse
import os

import numpy as np
import scipy.sparse as sp

from . import cusparse, host
from .sparse import csr_matrix


def save_csr_txt(prefix, csr):
    """gen_and_save_alg1_txt.py:8-14 (int32 / int32 / float32, one number per line)."""
    if isinstance(csr, csr_matrix):
        indptr, indices, data = csr.indptr.cpu().numpy(), csr.indices.cpu().numpy(), csr.data.cpu().numpy()
    else:
        indptr, indices, data = csr.indptr, csr.indices, csr.data
    host.write_csr_txt(prefix, indptr.astype(np.int32, copy=False), indices.astype(np.int32, copy=False),
                       data.astype(np.float32, copy=False))


def gen_rand_csr(n, d, seed):
    """gen_and_save_alg1_txt.py:16-19 with numpy's RandomState in place of cupy's."""
    rs = np.random.RandomState(seed)
    M = sp.random(n, n, density=d, format="csr", dtype=np.float32, random_state=rs)
    M.sort_indices()
    return M


def case_tag(n, d, alg, chunk_fraction=None, include_cf_in_tag=False):
    cf = f"_cf{str(chunk_fraction).replace('.', 'p')}" if (include_cf_in_tag and alg == 3) else ""
    return f"n{n}_dens{str(d).replace('.', 'p')}_alg{alg}{cf}"


def run_once(n, d, outdir, alg=1, chunk_fraction=0.2, seed=0, include_cf_in_tag=False):
    os.makedirs(outdir, exist_ok=True)
    A = gen_rand_csr(n, d, seed)
    B = gen_rand_csr(n, d, seed + 1)
    C = cusparse.spgemm(csr_matrix(A), csr_matrix(B), alg=alg, chunk_fraction=chunk_fraction)
    tag = case_tag(n, d, alg, chunk_fraction, include_cf_in_tag)
    save_csr_txt(os.path.join(outdir, f"A_{tag}"), A)
    save_csr_txt(os.path.join(outdir, f"B_{tag}"), B)
    save_csr_txt(os.path.join(outdir, f"C_py_{tag}"), C)
    extra = f", chunk_fraction={chunk_fraction}" if alg == 3 else ""
    print(f"[PY] saved A/B/C txt to {outdir} ({tag}{extra})")
    return tag


def main(argv=None, alg=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--sizes", nargs="+", type=int, default=[32, 64, 128, 256, 512, 1024])
    ap.add_argument("--densities", nargs="+", type=float, default=[0.01, 0.1, 0.3, 0.5])
    ap.add_argument("--outdir", default=None)
    ap.add_argument("--seed", type=int, default=123)
    ap.add_argument("--alg", type=int, default=alg or 1, choices=[1, 2, 3])
    ap.add_argument("--chunk-fraction", type=float, default=0.2, help="ALG3 chunk_fraction in (0,1], default=0.2")
    ap.add_argument("--include-cf-in-tag", action="store_true")
    args = ap.parse_args(argv)
    if not (0.0 < args.chunk_fraction <= 1.0):
        raise SystemExit(f"chunk_fraction must be in (0,1], got {args.chunk_fraction}")
    outdir = args.outdir or f"dump_alg{args.alg}_txt"
    for n in args.sizes:
        for d in args.densities:
            run_once(n, d, outdir, alg=args.alg, chunk_fraction=args.chunk_fraction, seed=args.seed,
                     include_cf_in_tag=args.include_cf_in_tag)


if __name__ == "__main__":
    main()
