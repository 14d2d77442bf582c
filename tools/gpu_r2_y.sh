#!/bin/bash
# round 2, GPU call Y: conversions (sum_duplicates fast path) parity
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_convert.py tests/test_gpu_compat.py -m gpu -x -q ) > gpurun_out/y_pytest_convert.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/y_pytest_convert.txt
