#!/bin/bash
# round 2, GPU call G: single-pass kernel with the late ticket, window turns of the blocked heavy-row kernel
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests/test_gpu_single_pass.py tests/test_gpu_spgemm.py -m gpu -x -q ) > gpurun_out/g_pytest_spgemm.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/g_pytest_spgemm.txt
timeout 300 python tools/single_pass_sweep.py > gpurun_out/g_single_pass_sweep.jsonl 2> gpurun_out/g_single_pass_sweep.err
timeout 900 python tools/heavy_modes.py 16 18 20 --configs blocked_w1,blocked_w4,blocked_w8,blocked_w4_f64,blocked_w8_f64,blocked_w8_f96,blocked_w8_f128 > gpurun_out/g_heavy_modes.jsonl 2> gpurun_out/g_heavy_modes.err
REPS=3 timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_num_dense_cta" -s 2 -c 1 \
   -o gpurun_out/g_ncu_c3 -f python tools/profile_cases.py c3_1e-2 > gpurun_out/g_ncu_c3.log 2>&1
