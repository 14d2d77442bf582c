#!/bin/bash
# round 2, GPU call W: heaviest rows first in the ticketed heavy-row kernels: parity, A/B at N=1
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_spgemm.py tests/test_gpu_baseline_sizes.py -m gpu -x -q ) > gpurun_out/w_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/w_pytest.txt
for rep in 1 2; do
  B200SP_NO_HEAVY_FIRST=1 timeout 600 python tools/heavy_modes.py 16 18 20 --configs blocked > gpurun_out/w_heavy_off_$rep.jsonl 2> gpurun_out/w_heavy_off_$rep.err
  timeout 600 python tools/heavy_modes.py 16 18 20 --configs blocked > gpurun_out/w_heavy_on_$rep.jsonl 2> gpurun_out/w_heavy_on_$rep.err
done
