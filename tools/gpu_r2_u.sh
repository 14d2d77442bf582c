#!/bin/bash
# round 2, GPU call U: launch list of COO -> CSR (bucket path) on the C2 recipe
mkdir -p gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"^k_" -c 200 --csv --log-file gpurun_out/u_launches_convert.csv \
   python tools/convert_bench.py --reps 2 > gpurun_out/u_convert.json 2> gpurun_out/u_convert.err
