#!/usr/bin/env python3
"""Drop-in for the reference's cupy_cusparse/gen_and_save_alg2_txt.py (same arguments, same
files): PYGEN_SCRIPT=scripts/gen_and_save_alg2_txt.py bash run_all_alg2.sh"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from spmm_b200 import corpus  # noqa: E402

if __name__ == "__main__":
    corpus.main(alg=2)
