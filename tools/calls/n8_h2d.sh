#!/bin/bash
set -u
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2n_topo.txt 2>&1; lscpu | grep -E "Socket|NUMA|Model name|^CPU\(s\)" >> gpurun_out/r2n_topo.txt
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29611 tools/h2d_probe.py > gpurun_out/r2n_h2d.jsonl 2> gpurun_out/r2n_h2d.err; echo "rc=$?" >> gpurun_out/r2n_h2d.err
grep '^{' gpurun_out/r2n_h2d.jsonl; tail -n 12 gpurun_out/r2n_topo.txt
