#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 850 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus 8 --steps 20 --warmup 5 --verbose > gpurun_out/r2p_bench.json 2> gpurun_out/r2p_bench.err; echo "bench rc=$?" >> gpurun_out/r2p_bench.err
grep '^{' gpurun_out/r2p_bench.json | cut -c1-220; grep -v "^\[W\|^W1\|^$" gpurun_out/r2p_bench.err | tail -n 14
