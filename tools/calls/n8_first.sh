#!/bin/bash
# N=8 call: the driver's form of the bench at 8 GPUs (north_star block included).
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name,memory.total --format=csv > gpurun_out/r2d_gpus.txt 2>&1; nproc >> gpurun_out/r2d_gpus.txt; free -g >> gpurun_out/r2d_gpus.txt
timeout 850 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 8 --steps 20 --warmup 5 --verbose > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err; echo "bench rc=$?" >> gpurun_out/r2d_bench.err
cut -c1-300 gpurun_out/r2d_bench.json; grep -v "^\[W\|^W1\|^$" gpurun_out/r2d_bench.err | tail -25
