#!/bin/bash
# round 2, GPU call V (8 GPUs): the N=8 bench line as the driver launches it (+: the C5 job also without the hub split)
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/v_topo.txt 2>&1
( time timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus 8 --steps 10 --warmup 3 ) > gpurun_out/v_bench_n8.json 2> gpurun_out/v_bench_n8.err
echo "bench rc=$?" >> gpurun_out/v_bench_n8.err
