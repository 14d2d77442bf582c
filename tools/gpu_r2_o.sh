#!/bin/bash
# round 2, GPU call O: k_num_warp_blocks (one warp per (row, narrow block)): parity, then timing against the blocked kernel
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_spgemm.py -m gpu -x -q -k "heavy_row_kernels_all_modes or hubs" ) > gpurun_out/o_pytest_heavy.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/o_pytest_heavy.txt
timeout 600 python tools/heavy_modes.py 16 18 --configs blocked,warp,warp_ww1280,warp_ww5120 > gpurun_out/o_heavy_modes.jsonl 2> gpurun_out/o_heavy_modes.err
B200SP_HEAVY=warp REPS=2 timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_num_warp_blocks" -s 1 -c 1 \
   -o gpurun_out/o_ncu_rmat18_warp -f python tools/profile_cases.py rmat18 > gpurun_out/o_ncu_rmat18.log 2>&1
