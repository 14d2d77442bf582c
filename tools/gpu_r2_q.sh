#!/bin/bash
# round 2, GPU call Q: whole-warp emit of well-filled bitmap words in the heavy-row sweeps: parity, A/B against the v19 library
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_spgemm.py -m gpu -x -q -k "heavy or hubs or rmat16 or window or all_bins or streamed" ) > gpurun_out/q_pytest_heavy.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/q_pytest_heavy.txt
for rep in 1 2; do
  B200SP_LIB=$PWD/tools/_build/libb200sp_v19.so timeout 600 python tools/heavy_modes.py 16 18 20 --configs blocked > gpurun_out/q_heavy_v19_$rep.jsonl 2> gpurun_out/q_heavy_v19_$rep.err
  timeout 600 python tools/heavy_modes.py 16 18 20 --configs blocked > gpurun_out/q_heavy_new_$rep.jsonl 2> gpurun_out/q_heavy_new_$rep.err
done
