#!/bin/bash
# N=8: final bench line (driver's form) + two unit-size variants of the config-2 step without the long legs.
set -u
mkdir -p gpurun_out
timeout 850 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus 8 --steps 20 --warmup 5 --verbose > gpurun_out/r2g_bench.json 2> gpurun_out/r2g_bench.err; echo "bench rc=$?" >> gpurun_out/r2g_bench.err
for U in 131072 65536; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29572 bench.py --gpus 8 --steps 50 --warmup 5 --unit-bytes $U --no-north-star --no-parity --no-e2e --no-extras > gpurun_out/r2g_bench_u$U.json 2> gpurun_out/r2g_bench_u$U.err; echo "rc=$?" >> gpurun_out/r2g_bench_u$U.err
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29573 bench.py --gpus 8 --steps 50 --warmup 5 --no-north-star --no-parity --no-e2e --no-extras > gpurun_out/r2g_bench_u262144.json 2> gpurun_out/r2g_bench_u262144.err
cut -c1-260 gpurun_out/r2g_bench.json; grep -v "^\[W\|^W1\|^$" gpurun_out/r2g_bench.err | tail -n 14; for f in gpurun_out/r2g_bench_u*.json; do echo $f; cut -c1-200 $f; done
