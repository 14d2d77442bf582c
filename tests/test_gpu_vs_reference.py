"""Our CLI drop-in against the reference's own cuSPARSE drivers on the same txt-CSR files.

oracle/_ref/spgemm_from_txt_alg{1,2,3} are the UNMODIFIED reference drivers
(cupy_cusparse/spgemm_from_txt_alg*.cu) compiled where they lie by oracle/Makefile; they
travel to the GPU box as prebuilt binaries.  This is the run_all_alg*.sh loop
(run_all_alg1.sh:43-74) with our executable on one side and the reference's on the other,
judged with compare_csrs_txt.py's criterion (:26-47): indptr and indices identical; values
identical up to summation order (cuSPARSE's order is unspecified, so a relative tolerance of
1e-6 of the row's absolute product sum replaces bitwise equality there).  The same outputs are
also checked bitwise against the oracle.
"""
import os
import subprocess

import numpy as np
import pytest

import helpers

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OURS = os.path.join(ROOT, "spmm_b200", "bin")
REF = os.path.join(ROOT, "oracle", "_ref")


def _save(prefix, M):  # gen_and_save_alg1_txt.py:8-14
    np.savetxt(f"{prefix}_indptr.txt", M.indptr, fmt="%d")
    np.savetxt(f"{prefix}_indices.txt", M.indices, fmt="%d")
    np.savetxt(f"{prefix}_data.txt", M.data, fmt="%.9g")


def _load(prefix):
    return (np.loadtxt(f"{prefix}_indptr.txt", dtype=np.int64, ndmin=1).astype(np.int32),
            np.loadtxt(f"{prefix}_indices.txt", dtype=np.int64, ndmin=1).astype(np.int32),
            np.loadtxt(f"{prefix}_data.txt", dtype=np.float64, ndmin=1).astype(np.float32))


def _run(exe, *args, env=None):
    r = subprocess.run([exe, *args], capture_output=True, text=True, env={**os.environ, **(env or {})})
    assert r.returncode == 0, f"{exe} failed: {r.stderr}"
    return r.stdout


@pytest.mark.parametrize("alg", [1, 2, 3])
@pytest.mark.parametrize("n,d", [(32, 0.1), (128, 0.3), (512, 0.1), (1024, 0.01), (1024, 0.3)])
def test_cli_matches_reference_driver(tmp_path, alg, n, d):
    ref_exe = os.path.join(REF, f"spgemm_from_txt_alg{alg}")
    if not os.path.exists(ref_exe):
        pytest.skip("oracle/_ref not built (reference checkout was absent at build time)")
    A, B = helpers.grid_case(n, d)
    tag = f"n{n}_dens{str(d).replace('.', 'p')}_alg{alg}"   # naming of gen_and_save_alg1_txt.py:26-33
    pa, pb = str(tmp_path / f"A_{tag}"), str(tmp_path / f"B_{tag}")
    _save(pa, A)
    _save(pb, B)
    c_ours, c_ref = str(tmp_path / f"C_ours_{tag}"), str(tmp_path / f"C_cu_{tag}")
    out = _run(os.path.join(OURS, f"spgemm_from_txt_alg{alg}"), pa, pb, c_ours)
    assert out.startswith(f"[C++] ALG{alg} wrote {c_ours}_*") and f"rows={n}" in out
    out_ref = _run(ref_exe, pa, pb, c_ref)
    # the reference's alg2 driver prints "ALG3" in its banner (spgemm_from_txt_alg2.cu:228), so only the prefix is checked
    assert out_ref.startswith("[C++] ALG") and f"rows={n}" in out_ref
    op, oj, ox = _load(c_ours)
    rp, rj, rx = _load(c_ref)
    assert np.array_equal(op, rp), "indptr differs from the reference driver"
    assert np.array_equal(oj, rj), "indices differ from the reference driver"
    absA, absB = abs(A), abs(B)
    bound = (absA @ absB).tocsr()
    bound.sort_indices()
    assert np.all(np.abs(ox - rx) <= 1e-6 * bound.data + 1e-30), "values differ beyond summation-order rounding"
    helpers.assert_csr_bitwise((op, oj, ox), helpers.oracle_spgemm(A, B), "cli vs oracle")


def test_cli_chunk_fraction_argument_and_env(tmp_path):
    A, B = helpers.grid_case(256, 0.1)
    pa, pb = str(tmp_path / "A"), str(tmp_path / "B")
    _save(pa, A)
    _save(pb, B)
    exe = os.path.join(OURS, "spgemm_from_txt_alg3")
    o1 = _run(exe, pa, pb, str(tmp_path / "C1"), "0.5")
    assert "chunk_fraction=0.5" in o1
    o2 = _run(exe, pa, pb, str(tmp_path / "C2"), env={"CHUNK_FRACTION": "0.1"})
    assert "chunk_fraction=0.1" in o2
    for f in ("indptr", "indices", "data"):   # twice-run cmp: deterministic/test_same.sh
        assert open(tmp_path / f"C1_{f}.txt").read() == open(tmp_path / f"C2_{f}.txt").read()


def test_reference_driver_keeps_zero_sums(tmp_path):
    """cuSPARSE (the reference's GPU path) keeps entries whose sum is exactly zero - the
    structural pattern - where scipy drops them; the product follows cuSPARSE."""
    ref_exe = os.path.join(REF, "spgemm_from_txt_alg1")
    if not os.path.exists(ref_exe):
        pytest.skip("oracle/_ref not built (reference checkout was absent at build time)")
    import scipy.sparse as sp
    A = sp.csr_matrix((np.array([1, 1, 0, 3], np.float32), np.array([0, 1, 0, 1]), np.array([0, 2, 4])), shape=(2, 2))
    B = sp.csr_matrix((np.array([2, 1, -2, 1], np.float32), np.array([0, 1, 0, 1]), np.array([0, 2, 4])), shape=(2, 2))
    pa, pb = str(tmp_path / "A"), str(tmp_path / "B")
    _save(pa, A)
    _save(pb, B)
    _run(ref_exe, pa, pb, str(tmp_path / "C_cu"))
    _run(os.path.join(OURS, "spgemm_from_txt_alg1"), pa, pb, str(tmp_path / "C_ours"))
    rp, rj, rx = _load(str(tmp_path / "C_cu"))
    op, oj, ox = _load(str(tmp_path / "C_ours"))
    assert rp.tolist() == op.tolist() == [0, 2, 4] and rj.tolist() == oj.tolist() == [0, 1, 0, 1]
    assert rx.tolist() == ox.tolist() == [0.0, 2.0, -6.0, 3.0]
    assert (A @ B).nnz == 3     # scipy stores only the non-zero sums
