#!/bin/bash
# round 2: the bench line on 2 GPUs (NCCL gather of config5), both arms as the driver launches them
set -u
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/g_bench_2gpu.json 2> gpurun_out/g_bench_2gpu.err; echo "bench 2gpu rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/g_ref_2gpu.json 2> gpurun_out/g_ref_2gpu.err; echo "ref rc=$?"
tail -c 1500 gpurun_out/g_bench_2gpu.json; tail -c 600 gpurun_out/g_bench_2gpu.err; tail -c 400 gpurun_out/g_ref_2gpu.json
