#!/bin/bash
# round 2, GPU call R: b200sp_spgemm_single_pass_direct: parity, then direct against estimate + single pass in one process
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests/test_gpu_single_pass.py tests/test_gpu_spgemm.py -m gpu -x -q ) > gpurun_out/r_pytest_spgemm.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r_pytest_spgemm.txt
for rep in 1 2; do
  B200SP_SINGLE_PASS_DIRECT=0 timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu --no-extra --no-c5 > gpurun_out/r_bench_estimate_$rep.json 2> gpurun_out/r_bench_estimate_$rep.err
  timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu --no-extra --no-c5 > gpurun_out/r_bench_direct_$rep.json 2> gpurun_out/r_bench_direct_$rep.err
done
