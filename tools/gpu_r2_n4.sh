#!/bin/bash
# round 2: the N=4 bench line as the driver launches it
mkdir -p gpurun_out
( time timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29514 \
    bench.py --gpus 4 --steps 10 --warmup 3 ) > gpurun_out/n4_bench.json 2> gpurun_out/n4_bench.err
echo "bench rc=$?" >> gpurun_out/n4_bench.err
