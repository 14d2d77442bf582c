#!/bin/bash
# round 2, GPU call AA: final build of round 2 (v22: prefetch depth 12 where three CTAs fit, hub split): whole suite, smoke, bench + reference arm, launch list, ncu --set full
mkdir -p gpurun_out
( time timeout 1800 python -m pytest tests -m gpu -x -q ) > gpurun_out/aa_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/aa_pytest.txt
( timeout 300 python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/aa_smoke.txt 2>&1
( time timeout 600 python bench.py --impl reference ) > gpurun_out/aa_bench_ref.json 2> gpurun_out/aa_bench_ref.err
( time timeout 900 python bench.py ) > gpurun_out/aa_bench_n1.json 2> gpurun_out/aa_bench_n1.err
echo "bench rc=$?" >> gpurun_out/aa_bench_n1.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/aa_launches.csv \
   python bench.py --steps 2 --warmup 1 --no-cpu --no-extra --no-c5 > gpurun_out/aa_ncu_launches.log 2>&1
REPS=3 timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_num_dense_cta|k_pack_b" -s 4 -c 2 \
   -o gpurun_out/aa_ncu_c3 -f python tools/profile_cases.py c3_1e-2 > gpurun_out/aa_ncu_c3.log 2>&1
