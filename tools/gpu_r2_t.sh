#!/bin/bash
# round 2, GPU call T: COO -> CSR by row buckets: parity (both paths), timing on the C2 recipe against the radix path
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_convert.py -m gpu -x -q ) > gpurun_out/t_pytest_convert.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/t_pytest_convert.txt
B200SP_COO_PATH=radix timeout 300 python tools/convert_bench.py > gpurun_out/t_convert_radix.json 2> gpurun_out/t_convert_radix.err
timeout 300 python tools/convert_bench.py > gpurun_out/t_convert_bucket.json 2> gpurun_out/t_convert_bucket.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"^k_" -c 60 --csv --log-file gpurun_out/t_launches_convert.csv \
   python tools/convert_bench.py --reps 2 > /dev/null 2> gpurun_out/t_ncu_convert.err
