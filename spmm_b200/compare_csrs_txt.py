"""Compare two txt-CSR matrices: same command line, messages and exit code as the reference's
cupy_cusparse/compare_csrs_txt.py (rows/nnz, indptr, indices and data must be identical; data
is compared as float32 after parsing, :26-47), reading the files with the library's fast parser
instead of numpy.loadtxt.

    python -m spmm_b200.compare_csrs_txt <py_prefix> <cu_prefix>
"""
import argparse

import numpy as np

from . import host


def load(prefix):
    indptr, indices, data = host.read_csr_txt(prefix)
    return indptr.size - 1, indices.size, indptr, indices, data


def main(py_prefix, cu_prefix):
    r1, n1, p1, i1, d1 = load(py_prefix)
    r2, n2, p2, i2, d2 = load(cu_prefix)
    ok = True
    if r1 != r2 or n1 != n2:
        print(f"rows/nnz mismatch: py=({r1},{n1}) cu=({r2},{n2})")
        ok = False
    if not np.array_equal(p1, p2):
        print("indptr mismatch")
        ok = False
    if not np.array_equal(i1, i2):
        print("indices mismatch")
        ok = False
    if not np.array_equal(d1, d2):
        print("data mismatch")
        ok = False
    print("EQUAL" if ok else "NOT EQUAL")
    return 0 if ok else 1


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("py_prefix")
    ap.add_argument("cu_prefix")
    a = ap.parse_args()
    raise SystemExit(main(a.py_prefix, a.cu_prefix))
