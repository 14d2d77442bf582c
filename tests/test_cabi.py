"""CPU-only checks of the drop-in boundary: the C-ABI library loads and exports every symbol
include/b200sp.h declares, the txt-CSR reader/writer and the CLI keep the reference's format
and exit codes, and the product fails loudly (no CPU fallback) when there is no CUDA device.
No compute call is made here."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

import helpers

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200sp.h")
CLI = os.path.join(ROOT, "spmm_b200", "bin", "b200sp_spgemm_from_txt")


def _declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200sp_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported_and_bound():
    from spmm_b200 import _lib
    lib = _lib.load()
    names = _declared_symbols()
    assert len(names) >= 20
    bound = {n for n, _, _ in _lib.SYMBOLS}
    for n in names:
        assert hasattr(lib, n), f"{n} declared in b200sp.h but not exported by libb200sp.so"
        assert n in bound, f"{n} declared in b200sp.h but not bound in spmm_b200/_lib.py"
    assert bound <= set(names), f"bound but undeclared: {bound - set(names)}"
    assert lib.b200sp_version() == 200
    assert lib.b200sp_status_string(-3) == b"out of memory"


def test_stats_struct_matches_header():
    from spmm_b200 import _lib
    # 5 int64 + 2 int32 + 2 * 16 int64 + int64 + 2 float
    assert ctypes.sizeof(_lib.Stats) == 5 * 8 + 8 + 2 * 16 * 8 + 8 + 8


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from spmm_b200 import _lib
    import spmm_b200
    hv = ctypes.c_void_p()
    assert _lib.load().b200sp_create(ctypes.byref(hv), 0) == _lib.ERR_CUDA
    with pytest.raises(RuntimeError):
        spmm_b200.csr_matrix(helpers.random_csr(8, 8, 0.5, 1))
    with pytest.raises(RuntimeError):
        spmm_b200.host.spgemm_host(helpers.random_csr(8, 8, 0.5, 1), helpers.random_csr(8, 8, 0.5, 2), (8, 8, 8))


def test_product_never_imports_oracle():
    """The oracle is test infrastructure: nothing shipped may import, link or dlopen it."""
    pat = re.compile(r"(import\s+oracle|from\s+oracle|liboracle|oracle/|oracle\.)")
    for dirpath, _, files in os.walk(os.path.join(ROOT, "spmm_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")) or f == "Makefile":
                txt = open(os.path.join(dirpath, f)).read()
                assert not pat.search(txt), f"{f} references the oracle"


# ---- txt-CSR format (spgemm_from_txt_alg1.cu:19-78, gen_and_save_alg1_txt.py:8-14) --------

def _savetxt_ref(prefix, M):
    """What the reference's generator writes (gen_and_save_alg1_txt.py:12-14)."""
    np.savetxt(f"{prefix}_indptr.txt", M.indptr, fmt="%d")
    np.savetxt(f"{prefix}_indices.txt", M.indices, fmt="%d")
    np.savetxt(f"{prefix}_data.txt", M.data, fmt="%.9g")


def test_txt_csr_roundtrip(tmp_path):
    from spmm_b200 import host
    M = helpers.random_csr(57, 33, 0.2, 9)
    M.data[:4] = np.array([1e-30, -3.4028235e38, 1.17549435e-38, 0.1], dtype=np.float32)
    pre = str(tmp_path / "A_n57")
    _savetxt_ref(pre, M)
    p, j, x = host.read_csr_txt(pre)
    assert np.array_equal(p, M.indptr) and np.array_equal(j, M.indices)
    assert np.array_equal(x.view(np.uint32), M.data.view(np.uint32))  # %.9g round-trips fp32
    out = str(tmp_path / "C_out")
    host.write_csr_txt(out, p, j, x)
    # the writer's files load back identically with numpy (compare_csrs_txt.py:26-47 semantics)
    assert np.array_equal(np.loadtxt(out + "_indptr.txt", dtype=np.int64), M.indptr)
    assert np.array_equal(np.loadtxt(out + "_indices.txt", dtype=np.int64), M.indices)
    assert np.array_equal(np.loadtxt(out + "_data.txt", dtype=np.float64).astype(np.float32), M.data)
    p2, j2, x2 = host.read_csr_txt(out)
    assert np.array_equal(x2.view(np.uint32), M.data.view(np.uint32))


def test_txt_csr_empty_and_errors(tmp_path):
    from spmm_b200 import host
    pre = str(tmp_path / "E")
    host.write_csr_txt(pre, np.zeros(4, np.int32), np.zeros(0, np.int32), np.zeros(0, np.float32))
    p, j, x = host.read_csr_txt(pre)
    assert p.tolist() == [0, 0, 0, 0] and j.size == 0 and x.size == 0
    with pytest.raises(OSError):
        host.read_csr_txt(str(tmp_path / "missing"))
    bad = str(tmp_path / "bad")
    open(bad + "_indptr.txt", "w").write("0\n2\n")
    open(bad + "_indices.txt", "w").write("0\nxyz\n")
    open(bad + "_data.txt", "w").write("1.0\n2.0\n")
    with pytest.raises((ValueError, OSError)):
        host.read_csr_txt(bad)


def test_cli_usage_and_validation(tmp_path):
    """Exit codes of the reference drivers: 2 usage (alg1.cu:118-121), 1 on bad input."""
    assert os.path.exists(CLI), "build the CLI first (__graft_entry__.build())"
    r = subprocess.run([CLI], capture_output=True, text=True)
    assert r.returncode == 2 and "A_prefix B_prefix C_prefix" in r.stderr + r.stdout
    r = subprocess.run([CLI, str(tmp_path / "noA"), str(tmp_path / "noB"), str(tmp_path / "C")],
                       capture_output=True, text=True)
    assert r.returncode == 1 and "noA_indptr.txt" in r.stderr
    r = subprocess.run([os.path.join(os.path.dirname(CLI), "spgemm_from_txt_alg3"), "a", "b", "c", "1.5"],
                       capture_output=True, text=True)
    assert r.returncode == 1 and "chunk_fraction" in r.stderr
    # out-of-range column index: validate_csr_indices (alg1.cu:78-99) -> exit 1 before any GPU work
    M = helpers.random_csr(6, 6, 0.5, 3)
    pa, pb = str(tmp_path / "A"), str(tmp_path / "B")
    _savetxt_ref(pa, M)
    M2 = M.copy()
    M2.indices[0] = 99
    _savetxt_ref(pb, M2)
    r = subprocess.run([CLI, pa, pb, str(tmp_path / "C")], capture_output=True, text=True)
    assert r.returncode == 1
    for k in (1, 2, 3):
        assert os.path.islink(os.path.join(os.path.dirname(CLI), f"spgemm_from_txt_alg{k}"))
