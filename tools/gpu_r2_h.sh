#!/bin/bash
# round 2, GPU call H: build v17 (single-pass SpGEMM, ticket at row start): whole suite, bench line + reference arm, launch list, ncu --set full
mkdir -p gpurun_out
( time timeout 1800 python -m pytest tests -m gpu -x -q ) > gpurun_out/h_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/h_pytest.txt
( time timeout 900 python bench.py ) > gpurun_out/h_bench_n1.json 2> gpurun_out/h_bench_n1.err
echo "bench rc=$?" >> gpurun_out/h_bench_n1.err
( time timeout 600 python bench.py --impl reference ) > gpurun_out/h_bench_ref.json 2> gpurun_out/h_bench_ref.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/h_launches.csv \
   python bench.py --steps 2 --warmup 1 --no-cpu --no-extra --no-c5 > gpurun_out/h_ncu_launches.log 2>&1
REPS=3 timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_num_dense_cta" -s 2 -c 1 \
   -o gpurun_out/h_ncu_c3 -f python tools/profile_cases.py c3_1e-2 > gpurun_out/h_ncu_c3.log 2>&1
