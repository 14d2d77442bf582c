#!/bin/bash
# round 2, GPU call D: the allocation-free SpMV entry (alone in a fresh process, then at the end of the whole suite),
# heavy-row kernel modes on R-MAT, SpMV sweep, ncu --set full of the SpMV kernels on C2
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q ) > gpurun_out/d_pytest_zz_alone.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/d_pytest_zz_alone.txt
( time timeout 1800 python -m pytest tests -m gpu -x -q ) > gpurun_out/d_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/d_pytest.txt
timeout 900 python tools/heavy_modes.py 16 18 20 > gpurun_out/d_heavy_modes.jsonl 2> gpurun_out/d_heavy_modes.err
timeout 300 python tools/spmv_sweep.py > gpurun_out/d_spmv_sweep.jsonl 2> gpurun_out/d_spmv_sweep.err
for k in vector merge; do
  B200SP_SPMV_KERNEL=$k REPS=3 python tools/profile_cases.py spmv > /dev/null 2>&1 && \
  B200SP_SPMV_KERNEL=$k REPS=3 timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_spmv_(vector|merge)" -s 1 -c 1 \
     -o gpurun_out/d_ncu_spmv_$k -f python tools/profile_cases.py spmv > gpurun_out/d_ncu_spmv_$k.log 2>&1
done
REPS=2 python tools/profile_cases.py rmat18 > /dev/null 2>&1 && \
REPS=2 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/d_launches_rmat18.csv \
   python tools/profile_cases.py rmat18 > gpurun_out/d_ncu_rmat18_launches.log 2>&1
ls -la gpurun_out > gpurun_out/d_ls.txt
