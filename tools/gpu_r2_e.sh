#!/bin/bash
# round 2, GPU call E: single-pass SpGEMM (tests, density sweep, bench line), heavy rows back on the blocked kernel
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_single_pass.py tests/test_gpu_spgemm.py -m gpu -x -q ) > gpurun_out/e_pytest_single.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/e_pytest_single.txt
timeout 600 python tools/single_pass_sweep.py > gpurun_out/e_single_pass_sweep.jsonl 2> gpurun_out/e_single_pass_sweep.err
( time timeout 900 python bench.py ) > gpurun_out/e_bench_n1.json 2> gpurun_out/e_bench_n1.err
echo "bench rc=$?" >> gpurun_out/e_bench_n1.err
( time timeout 1800 python -m pytest tests -m gpu -x -q ) > gpurun_out/e_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/e_pytest.txt
