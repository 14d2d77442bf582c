#!/bin/bash
# round 2, GPU call AD: blocked heavy-row kernel variants built from the same sources (window threshold 16 / 48, prefetch depth 8 / 24) against the default, one call
mkdir -p gpurun_out
for v in default flat16 flat48 bd8 bd24 default; do
  if [ $v = default ]; then unset B200SP_LIB; else export B200SP_LIB=$PWD/tools/_build/libb200sp_$v.so; fi
  timeout 600 python tools/heavy_modes.py 18 20 --configs blocked >> gpurun_out/ad_heavy_$v.jsonl 2>> gpurun_out/ad_heavy_$v.err
done
