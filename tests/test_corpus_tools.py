"""The txt-CSR corpus tools: naming and the comparison script on the
CPU; the generate -> compute -> compare loop of run_all_alg*.sh on the GPU."""
import os
import subprocess
import sys

import numpy as np
import pytest

import helpers

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CMP = os.path.join(ROOT, "scripts", "compare_csrs_txt.py")


def test_case_tag_matches_reference_naming():
    from spmm_b200 import corpus
    # gen_and_save_alg1_txt.py:29, gen_and_save_alg3_txt.py:30-31
    assert corpus.case_tag(1024, 0.01, 1) == "n1024_dens0p01_alg1"
    assert corpus.case_tag(32, 0.5, 2) == "n32_dens0p5_alg2"
    assert corpus.case_tag(64, 0.3, 3, 0.2, include_cf_in_tag=True) == "n64_dens0p3_alg3_cf0p2"
    assert corpus.case_tag(64, 0.3, 3, 0.2) == "n64_dens0p3_alg3"
    A = corpus.gen_rand_csr(64, 0.1, 123)
    B = helpers.random_csr(64, 64, 0.1, 123)
    assert np.array_equal(A.indices, B.indices) and np.array_equal(A.data, B.data) and A.has_sorted_indices


def test_compare_script_messages_and_exit_codes(tmp_path):
    from spmm_b200 import corpus
    A = helpers.random_csr(40, 40, 0.2, 1)
    corpus.save_csr_txt(str(tmp_path / "a"), A)
    corpus.save_csr_txt(str(tmp_path / "same"), A)
    B = A.copy()
    B.data[7] = np.nextafter(B.data[7], np.float32(10))
    corpus.save_csr_txt(str(tmp_path / "val"), B)
    C = helpers.random_csr(40, 40, 0.2, 2)
    corpus.save_csr_txt(str(tmp_path / "other"), C)

    def run(x, y):
        r = subprocess.run([sys.executable, CMP, str(tmp_path / x), str(tmp_path / y)], capture_output=True, text=True)
        return r.returncode, r.stdout.strip().splitlines()
    assert run("a", "same") == (0, ["EQUAL"])
    assert run("a", "val") == (1, ["data mismatch", "NOT EQUAL"])       # one ulp is a mismatch (compare_csrs_txt.py:43-45)
    rc, out = run("a", "other")
    assert rc == 1 and out[-1] == "NOT EQUAL" and any("mismatch" in line for line in out[:-1])
    r = subprocess.run([sys.executable, CMP, str(tmp_path / "a")], capture_output=True, text=True)
    assert r.returncode == 2                                            # argparse usage error


@pytest.mark.gpu
@pytest.mark.parametrize("alg", [1, 2, 3])
def test_run_all_loop(tmp_path, alg):
    """run_all_alg1.sh:26-74: C written by the Python operator == C written by the CLI, bit for
    bit in the text files, for every case; and both equal the oracle."""
    out = str(tmp_path / f"dump_alg{alg}_txt")
    env = {**os.environ, "ALG": str(alg), "SIZES": "32 128 256", "DENSITIES": "0.01 0.3", "PYGEN_PY": sys.executable,
           "CMPPY_PY": sys.executable, "CHUNK_FRACTION": "0.3"}
    r = subprocess.run(["bash", os.path.join(ROOT, "scripts", "run_all_alg.sh"), out], capture_output=True, text=True, env=env)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "6 PASS / 0 FAIL / 6 TOTAL" in r.stdout
    report = open(os.path.join(out, f"report_alg{alg}.txt")).read().splitlines()
    assert len(report) == 6 and all(line.startswith("[PASS] n") for line in report)
    from spmm_b200 import host
    tag = f"n256_dens0p3_alg{alg}"
    A, B = helpers.grid_case(256, 0.3)
    got = host.read_csr_txt(os.path.join(out, f"C_cu_{tag}"))
    want = helpers.oracle_spgemm(A, B)
    # text carries 9 significant digits: float32 round-trips exactly
    helpers.assert_csr_bitwise(got, want, "txt corpus vs oracle")


@pytest.mark.gpu
def test_run_all_loop_reports_failure(tmp_path):
    out = str(tmp_path / "d")
    fake = tmp_path / "fake_exe.sh"
    fake.write_text("#!/usr/bin/env bash\nfor f in indptr indices data; do cp \"$1_$f.txt\" \"$3_$f.txt\"; done\n")  # writes A as C
    fake.chmod(0o755)
    env = {**os.environ, "ALG": "1", "SIZES": "32", "DENSITIES": "0.3", "PYGEN_PY": sys.executable, "CMPPY_PY": sys.executable,
           "CUEXE": str(fake)}
    r = subprocess.run(["bash", os.path.join(ROOT, "scripts", "run_all_alg.sh"), out], capture_output=True, text=True, env=env)
    assert r.returncode == 1 and "[FAIL] n32_dens0p3_alg1" in r.stdout and "0 PASS / 1 FAIL / 1 TOTAL" in r.stdout
