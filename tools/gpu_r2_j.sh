#!/bin/bash
# round 2, GPU call J: sort_indices fallback test, source-level ncu captures of the heavy-row kernel at R-MAT scale 18 and 20
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_convert.py -m gpu -x -q ) > gpurun_out/j_pytest_convert.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/j_pytest_convert.txt
REPS=2 timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_num_dense_pieces" -s 1 -c 1 \
   -o gpurun_out/j_ncu_rmat18 -f python tools/profile_cases.py rmat18 > gpurun_out/j_ncu_rmat18.log 2>&1
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"k_num_dense_pieces" -s 4 -c 1 \
   -o gpurun_out/j_ncu_rmat20 -f python tools/heavy_modes.py 20 --configs blocked --reps 1 > gpurun_out/j_ncu_rmat20.log 2>&1
