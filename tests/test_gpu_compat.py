"""The reference's own Python scripts, UNCHANGED, on this library: launched through
`python -m spmm_b200.compat` (the cupy / cupyx alias layer) from the copies that oracle/Makefile
stages next to the compiled reference drivers (oracle/_ref/scripts, git-ignored; /root/reference
does not exist on the GPU box).  Skipped when the staged scripts are absent."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCRIPTS = os.path.join(ROOT, "oracle", "_ref", "scripts")

needs_scripts = pytest.mark.skipif(not os.path.isdir(os.path.join(SCRIPTS, "SpGEMM_alg_comparison")),
                                   reason="reference scripts not staged (run `make -C oracle` where /root/reference exists)")


def launch(script, *args, cwd=None, stub_plots=False, timeout=600):
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""),
               PYTORCH_CUDA_ALLOC_CONF="expandable_segments:True")
    cmd = [sys.executable, "-m", "spmm_b200.compat"] + (["--stub-matplotlib"] if stub_plots else []) + \
          [os.path.join(SCRIPTS, script)] + [str(a) for a in args]
    r = subprocess.run(cmd, capture_output=True, text=True, cwd=cwd or ROOT, env=env, timeout=timeout)
    assert r.returncode == 0, f"{script} exited {r.returncode}\n{r.stdout[-2000:]}\n{r.stderr[-3000:]}"
    return r.stdout


@needs_scripts
def test_alg_comparison_profiler_runs_unchanged():
    """SpGEMM_alg_comparison/profiler.py --size 1024 --density 0.01 --runs 3: the result table of
    profiler.py:224-230 with one line per alg."""
    out = launch("SpGEMM_alg_comparison/profiler.py", "--size", 1024, "--density", 0.01, "--runs", 3)
    assert "=== Results (alg comparison) ===" in out
    rows = [ln for ln in out.splitlines() if ln.startswith("A_csr @ B_csr (alg=")]
    assert len(rows) == 3
    for alg, ln in zip((1, 2, 3), rows):
        f = ln.split()
        assert f"(alg={alg})" in ln and float(f[4]) > 0.0 and "(1024, 1024)" in ln and "float32" in ln


@needs_scripts
def test_spgemm_vs_spmv_profiler_runs_unchanged():
    """SpGEMM_vs_SpMV/profiler.py: the {csr,csc,coo}^2 SpGEMM grid and the three (n,1) SpMV rows
    (profiler.py:485-501), CPU (scipy) section included."""
    out = launch("SpGEMM_vs_SpMV/profiler.py", "--m", 512, "--n", 512, "--p", 512, "--densityA", 0.02, "--densityB", 0.02,
                 "--runs", 2)
    gpu = out[out.index("=== Config (GPU/cupy) ==="):]
    assert sum(1 for ln in gpu.splitlines() if "(SpGEMM)" in ln and ln.startswith("A_")) == 9
    spmv = [ln for ln in gpu.splitlines() if "(SpMV, dense vec)" in ln and ln.startswith("A_")]
    assert len(spmv) == 3 and all("(512, 1)" in ln for ln in spmv)


@needs_scripts
def test_dense_vs_sparse_runs_unchanged():
    out = launch("dense_vs_sparseGEMM/main.py", "--size", 512, "--density", 0.001, 0.01, "--runs", 2)
    assert out.count("A @ B [sparse, inputs_on_gpu]") == 2 and out.count("A @ B [dense, inputs_on_gpu]") == 2
    assert "[SKIP]" not in out


@needs_scripts
@pytest.mark.parametrize("alg", [1, 2, 3])
def test_deterministic_scripts_twice_identical(alg, tmp_path):
    """deterministic/test_deterministic.sh:4-36: the script twice, `diff` of what it wrote."""
    f1, f2 = tmp_path / "file1.txt", tmp_path / "file2.txt"
    launch(f"deterministic/cupy_alg{alg}.py", "--out", f1, "--dtype", "float32")
    launch(f"deterministic/cupy_alg{alg}.py", "--out", f2, "--dtype", "float32")
    a, b = f1.read_text(), f2.read_text()
    assert len(a) > 1000 and a == b
    assert subprocess.run(["diff", "-q", str(f1), str(f2)]).returncode == 0


@needs_scripts
def test_numerical_error_study_runs_with_plot_stub(tmp_path):
    """numerical_error/error.py (ALG1 vs ALG3 max |diff| over a size x density grid, :18-29): every
    mode of this library gives identical bits, so every cell must read 0."""
    out = launch("numerical_error/error.py", cwd=str(tmp_path), stub_plots=True)
    assert "cell texts" in out
    cells = out[out.index("cell texts"):]
    assert cells.count("'0.00e+00'") == 9


@needs_scripts
def test_generator_script_feeds_the_cli(tmp_path):
    """cupy_cusparse/gen_and_save_alg3_txt.py (the PYGEN_SCRIPT of run_all_alg3.sh) writes A, B and
    C_py through the alias layer; the CLI driver reproduces C_py bit for bit (compare_csrs_txt.py)."""
    launch("cupy_cusparse/gen_and_save_alg3_txt.py", "--sizes", 64, "--densities", 0.1, "--outdir", tmp_path)
    names = sorted(p.name for p in tmp_path.iterdir())
    pre = [n[:-len("_indptr.txt")] for n in names if n.endswith("_indptr.txt")]
    a = next(p for p in pre if p.startswith("A_"))
    b = next(p for p in pre if p.startswith("B_"))
    c = next(p for p in pre if p.startswith("C_py"))
    exe = os.path.join(ROOT, "spmm_b200", "bin", "spgemm_from_txt_alg3")
    r = subprocess.run([exe, str(tmp_path / a), str(tmp_path / b), str(tmp_path / "C_cu")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    cmp_ = subprocess.run([sys.executable, os.path.join(SCRIPTS, "cupy_cusparse", "compare_csrs_txt.py"),
                           str(tmp_path / c), str(tmp_path / "C_cu")], capture_output=True, text=True)
    assert cmp_.returncode == 0 and "EQUAL" in cmp_.stdout and "NOT EQUAL" not in cmp_.stdout, cmp_.stdout + cmp_.stderr
