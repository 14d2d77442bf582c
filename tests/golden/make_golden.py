#!/usr/bin/env python3
"""Generates the golden vectors under tests/golden/ from the reference's CPU path.

The reference's CPU implementation of the hot path is `scipy.sparse` `A @ B` / `A @ x`
(SpGEMM_vs_SpMV/profiler.py:408,411); the reference's GPU side needs CuPy + a GPU and cannot
run in the build container, and the repo commits no golden outputs of its own (SURVEY.md
section 8c).  So the vectors are produced HERE by scipy (version recorded in the files), with
the reference's own input recipes:

  * grid cases: cupy_cusparse/gen_and_save_alg1_txt.py:16-24 (random CSR, seeds 123 / 124,
    sort_indices) on the size x density grid of cupy_cusparse/run_all_alg1.sh:5-7;
  * C1 / C3: SpGEMM_vs_SpMV/profiler.py:296-302 (default_rng(42), standard_normal values,
    eliminate_zeros; B built from the same seed, so B == A).

Outputs
  spgemm_small.npz   full A, B, C arrays of a few small cases (fp32 + fp64, rectangular, alpha)
  spmv_small.npz     A, x, y for SpMV cases
  convert_small.npz  COO -> canonical CSR (duplicates at most two-fold, so SciPy's bits are
                     well defined) and CSC -> CSR cases (scipy tocsr / sum_duplicates)
  checksums.json     sha256 of (indptr, indices, data bytes) of C for the whole 24-case grid, C1
                     and the light end of the C3 sweep - cases too big to commit as arrays

Run:  python tests/golden/make_golden.py      (rewrites the three files; deterministic)
"""
import hashlib
import json
import os
import sys

import numpy as np
import scipy
import scipy.sparse as sp

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import helpers  # noqa: E402  (tests/helpers.py: the input recipes)


def scipy_product(A, B, alpha=None):
    C = A @ B
    if alpha is not None:
        C = C * A.dtype.type(alpha)
    C = C.tocsr()
    C.sort_indices()
    return C


def sha(C):
    h = hashlib.sha256()
    h.update(np.ascontiguousarray(C.indptr, dtype=np.int32).tobytes())
    h.update(np.ascontiguousarray(C.indices, dtype=np.int32).tobytes())
    h.update(np.ascontiguousarray(C.data).tobytes())
    return h.hexdigest()


def small_cases():
    cases = {}
    for name, (n, d, dt) in {"grid_n32_d0p1_f32": (32, 0.1, np.float32), "grid_n64_d0p3_f32": (64, 0.3, np.float32),
                             "grid_n128_d0p01_f32": (128, 0.01, np.float32), "grid_n64_d0p1_f64": (64, 0.1, np.float64),
                             "grid_n256_d0p1_f32": (256, 0.1, np.float32)}.items():
        A, B = helpers.grid_case(n, d, dtype=dt)
        cases[name] = (A, B, None)
    A = helpers.random_csr(37, 91, 0.08, 5)
    B = helpers.random_csr(91, 53, 0.11, 6)
    cases["rect_37x91x53_f32"] = (A, B, None)
    cases["rect_alpha0p5_f32"] = (A, B, 0.5)
    A = helpers.profiler_matrix(300, 300, 0.02, seed=42)
    cases["profiler_300_A_at_A_f32"] = (A, A, None)
    return cases


def main():
    arrays = {}
    for name, (A, B, alpha) in small_cases().items():
        C = scipy_product(A, B, alpha)
        for tag, M in (("A", A), ("B", B), ("C", C)):
            arrays[f"{name}/{tag}p"] = M.indptr.astype(np.int32)
            arrays[f"{name}/{tag}j"] = M.indices.astype(np.int32)
            arrays[f"{name}/{tag}x"] = M.data
        arrays[f"{name}/shape"] = np.array([A.shape[0], A.shape[1], B.shape[1]], dtype=np.int64)
        arrays[f"{name}/alpha"] = np.array(1.0 if alpha is None else alpha)
    np.savez_compressed(os.path.join(HERE, "spgemm_small.npz"), **arrays)

    mv = {}
    rng = np.random.default_rng(11)
    for name, (m, n, d, dt) in {"mv_50x70_f32": (50, 70, 0.2, np.float32), "mv_200x200_f64": (200, 200, 0.05, np.float64),
                                "mv_1x400_f32": (1, 400, 0.9, np.float32)}.items():
        A = helpers.random_csr(m, n, d, 21, dtype=dt)
        x = rng.standard_normal(n).astype(dt)
        mv[f"{name}/Ap"], mv[f"{name}/Aj"], mv[f"{name}/Ax"] = A.indptr.astype(np.int32), A.indices.astype(np.int32), A.data
        mv[f"{name}/x"] = x
        mv[f"{name}/y"] = A @ x
        mv[f"{name}/shape"] = np.array(A.shape, dtype=np.int64)
    np.savez_compressed(os.path.join(HERE, "spmv_small.npz"), **mv)

    cv = {}
    for name, (m, n, k, seed, dt) in {"coo_pairs_40x30_f32": (40, 30, 300, 3, np.float32),
                                      "coo_pairs_1x64_f64": (1, 64, 40, 4, np.float64),
                                      "coo_pairs_257x3_f32": (257, 3, 500, 5, np.float32)}.items():
        row, col, val = helpers.coo_with_pairs(m, n, k, seed, dt)
        M = helpers.scipy_canonical_from_coo(row, col, val, (m, n))
        cv[f"{name}/row"], cv[f"{name}/col"], cv[f"{name}/val"] = row, col, val
        cv[f"{name}/Cp"], cv[f"{name}/Cj"], cv[f"{name}/Cx"] = M.indptr.astype(np.int32), M.indices.astype(np.int32), M.data
        cv[f"{name}/shape"] = np.array([m, n], dtype=np.int64)
    for name, (m, n, d, dt) in {"csc_50x70_f32": (50, 70, 0.2, np.float32), "csc_9x300_f64": (9, 300, 0.1, np.float64)}.items():
        M = sp.random(m, n, density=d, format="csc", dtype=dt, random_state=np.random.RandomState(8))
        R = M.tocsr()
        cv[f"{name}/ptr"], cv[f"{name}/idx"], cv[f"{name}/val"] = M.indptr.astype(np.int32), M.indices.astype(np.int32), M.data
        cv[f"{name}/Cp"], cv[f"{name}/Cj"], cv[f"{name}/Cx"] = R.indptr.astype(np.int32), R.indices.astype(np.int32), R.data
        cv[f"{name}/shape"] = np.array([m, n], dtype=np.int64)
    np.savez_compressed(os.path.join(HERE, "convert_small.npz"), **cv)

    sums = {"_scipy": scipy.__version__, "_numpy": np.__version__, "cases": {}}
    for n in helpers.GRID_SIZES:
        for d in helpers.GRID_DENSITIES:
            A, B = helpers.grid_case(n, d)
            C = scipy_product(A, B)
            sums["cases"][f"grid_n{n}_d{d:g}"] = {"nnzC": int(C.nnz), "sha256": sha(C)}
    A = helpers.profiler_matrix(4096, 4096, 1e-3)
    C = scipy_product(A, A)
    sums["cases"]["c1_4096_d1e-3"] = {"nnzA": int(A.nnz), "nnzC": int(C.nnz), "sha256": sha(C),
                                      "products": int(np.diff(A.indptr).astype(np.int64)[A.indices].sum())}
    for d in (1e-4, 1e-3):
        A = helpers.profiler_matrix(16384, 16384, d)
        C = scipy_product(A, A)
        sums["cases"][f"c3_16384_d{d:g}"] = {"nnzA": int(A.nnz), "nnzC": int(C.nnz), "sha256": sha(C),
                                             "products": int(np.diff(A.indptr).astype(np.int64)[A.indices].sum())}
    with open(os.path.join(HERE, "checksums.json"), "w") as f:
        json.dump(sums, f, indent=1, sort_keys=True)
    print("wrote", len(arrays), "arrays,", len(sums["cases"]), "checksums; scipy", scipy.__version__)


if __name__ == "__main__":
    main()
