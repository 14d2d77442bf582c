#!/bin/bash
# round 2, GPU call Z: sweep of the blocked kernel with compacted non-empty bitmap words: parity, A/B against the previous build
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_spgemm.py tests/test_gpu_baseline_sizes.py -m gpu -x -q ) > gpurun_out/z_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/z_pytest.txt
for rep in 1 2; do
  B200SP_LIB=$PWD/tools/_build/libb200sp_head.so timeout 600 python tools/heavy_modes.py 16 18 20 --configs blocked > gpurun_out/z_heavy_head_$rep.jsonl 2> gpurun_out/z_heavy_head_$rep.err
  timeout 600 python tools/heavy_modes.py 16 18 20 --configs blocked > gpurun_out/z_heavy_new_$rep.jsonl 2> gpurun_out/z_heavy_new_$rep.err
done
