#!/bin/bash
# round 2, GPU call AC: dense-kernel variants built from the same sources (prefetch depth 12 / 16, 256-thread CTAs) against the default, one call
mkdir -p gpurun_out
for v in default d12 d16 nt256 default; do
  if [ $v = default ]; then unset B200SP_LIB; else export B200SP_LIB=$PWD/tools/_build/libb200sp_$v.so; fi
  timeout 300 python tools/single_pass_sweep.py >> gpurun_out/ac_sweep_$v.jsonl 2>> gpurun_out/ac_sweep_$v.err
done
