#!/bin/bash
# round 2, last GPU call: the default bench at N=1 (short) and N=2 after the last edits of bench.py
mkdir -p gpurun_out
( timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu ) > gpurun_out/final_bench_n1.json 2> gpurun_out/final_bench_n1.err
echo "bench rc=$?" >> gpurun_out/final_bench_n1.err
( timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus 2 --steps 10 --warmup 3 ) > gpurun_out/final_bench_n2.json 2> gpurun_out/final_bench_n2.err
echo "bench rc=$?" >> gpurun_out/final_bench_n2.err
