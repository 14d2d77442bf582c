#!/bin/bash
# round 2, GPU call M: packed bank-dealt B for the dense kernel: parity tests, then A/B against the v18 library in one call
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests/test_gpu_single_pass.py tests/test_gpu_spgemm.py tests/test_gpu_baseline_sizes.py -m gpu -x -q ) > gpurun_out/m_pytest_spgemm.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/m_pytest_spgemm.txt
for rep in 1 2; do
  B200SP_LIB=$PWD/tools/_build/libb200sp_v18.so timeout 300 python tools/single_pass_sweep.py > gpurun_out/m_sweep_v18_$rep.jsonl 2> gpurun_out/m_sweep_v18_$rep.err
  timeout 300 python tools/single_pass_sweep.py > gpurun_out/m_sweep_new_$rep.jsonl 2> gpurun_out/m_sweep_new_$rep.err
done
REPS=3 timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_num_dense_cta|k_pack_b" -s 4 -c 2 \
   -o gpurun_out/m_ncu_c3 -f python tools/profile_cases.py c3_1e-2 > gpurun_out/m_ncu_c3.log 2>&1
