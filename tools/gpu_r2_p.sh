#!/bin/bash
# round 2, GPU call P: how fast is k_num_warp_blocks without the look-back wait (timing experiment, offsets wrong)
mkdir -p gpurun_out
timeout 600 python tools/heavy_modes.py 16 18 --configs blocked,warp_nowait,warp_nowait_ww5120 --reps 1 > gpurun_out/p_heavy_modes.jsonl 2> gpurun_out/p_heavy_modes.err
B200SP_HEAVY=warp B200SP_WB_NOWAIT=1 REPS=2 timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_num_warp_blocks" -s 1 -c 1 \
   -o gpurun_out/p_ncu_rmat18_warp -f python tools/profile_cases.py rmat18 > gpurun_out/p_ncu_rmat18.log 2>&1
