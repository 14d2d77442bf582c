#!/bin/bash
# round 2, GPU call B: GPU tests (rank kernel default, compat scripts), heavy-row kernel modes, SpMV sweep
mkdir -p gpurun_out
( time timeout 1800 python -m pytest tests -m gpu -q ) > gpurun_out/b_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/b_pytest.txt
timeout 900 python tools/heavy_modes.py 16 18 20 > gpurun_out/b_heavy_modes.jsonl 2> gpurun_out/b_heavy_modes.err
timeout 300 python tools/spmv_sweep.py > gpurun_out/b_spmv_sweep.jsonl 2> gpurun_out/b_spmv_sweep.err
