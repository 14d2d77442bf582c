#!/bin/bash
# round 2, GPU call A: full GPU test suite, N=1 bench line, ncu of the SpMV kernels on C2
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/a_gpu.txt
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/a_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/a_pytest.txt
( time timeout 900 python bench.py ) > gpurun_out/a_bench_n1.json 2> gpurun_out/a_bench_n1.err
echo "bench rc=$?" >> gpurun_out/a_bench_n1.err
for k in vector merge; do
  B200SP_SPMV_KERNEL=$k REPS=3 python tools/profile_cases.py spmv > /dev/null 2>&1 && \
  B200SP_SPMV_KERNEL=$k REPS=3 timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_spmv_(vector4|merge)" -s 1 -c 1 \
     -o gpurun_out/a_ncu_spmv_$k -f python tools/profile_cases.py spmv > gpurun_out/a_ncu_spmv_$k.log 2>&1
done
ls -la gpurun_out > gpurun_out/a_ls.txt
