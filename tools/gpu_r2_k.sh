#!/bin/bash
# round 2, GPU call K: branch-free emit of the dense kernel, CTA shapes of the blocked kernel, native row_products / handle pool
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests/test_gpu_single_pass.py tests/test_gpu_spgemm.py -m gpu -x -q ) > gpurun_out/k_pytest_spgemm.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/k_pytest_spgemm.txt
timeout 300 python tools/single_pass_sweep.py > gpurun_out/k_single_pass_sweep.jsonl 2> gpurun_out/k_single_pass_sweep.err
timeout 900 python tools/heavy_modes.py 16 18 20 --configs blocked,nt128_kb100,nt128_kb64,nt128_kb48,nt256_kb48,nt256_kb200 > gpurun_out/k_heavy_modes.jsonl 2> gpurun_out/k_heavy_modes.err
