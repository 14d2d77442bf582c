#!/bin/bash
# round 2, GPU call AB: hub rows split into (row, block) tasks: parity, A/B at N=1
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_spgemm.py tests/test_gpu_baseline_sizes.py -m gpu -x -q ) > gpurun_out/ab_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/ab_pytest.txt
for rep in 1 2; do
  B200SP_NO_HUB_SPLIT=1 timeout 600 python tools/heavy_modes.py 16 18 20 --configs blocked > gpurun_out/ab_heavy_off_$rep.jsonl 2> gpurun_out/ab_heavy_off_$rep.err
  timeout 600 python tools/heavy_modes.py 16 18 20 --configs blocked > gpurun_out/ab_heavy_on_$rep.jsonl 2> gpurun_out/ab_heavy_on_$rep.err
done
