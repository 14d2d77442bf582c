#!/usr/bin/env python3
"""Drop-in for the reference's cupy_cusparse/compare_csrs_txt.py (CMPPY_SCRIPT=...)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from spmm_b200 import compare_csrs_txt  # noqa: E402

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("py_prefix")
    ap.add_argument("cu_prefix")
    a = ap.parse_args()
    raise SystemExit(compare_csrs_txt.main(a.py_prefix, a.cu_prefix))
