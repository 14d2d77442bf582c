#!/bin/bash
# round 2, GPU call I (2 GPUs): the N=2 bench line as the driver launches it (weak-scaling C3 with B broadcast inside the step,
# e2e through multi.spgemm_row_panels, extra.c5_strong at scale 20, spmv_c2_multi)
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/i_topo.txt 2>&1
( time timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus 2 --steps 10 --warmup 3 ) > gpurun_out/i_bench_n2.json 2> gpurun_out/i_bench_n2.err
echo "bench rc=$?" >> gpurun_out/i_bench_n2.err
