#!/bin/bash
# N=4 call: bench --gpus 4 with a small north-star block (exercises that code path), multicast probe, ncu of a multicast launch.
set -u
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 4 --steps 20 --warmup 5 --verbose --north-star --north-star-layers 16 --north-star-size 268435456 --deadline 600 > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err; echo "bench rc=$?" >> gpurun_out/r2c_bench.err
timeout 400 python tools/mcast_probe.py --layers 4 --size 1073741824 --dsts 2,3 --quick > gpurun_out/r2c_mcast_probe.jsonl 2> gpurun_out/r2c_mcast_probe.err; echo "probe rc=$?" >> gpurun_out/r2c_mcast_probe.err
ONE="python tools/mcast_probe.py --layers 2 --size 536870912 --one"
timeout 120 $ONE > gpurun_out/r2c_one_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:copy_csum_kernel -s 1 -c 1 -o gpurun_out/r2c_mcast_prof $ONE > gpurun_out/r2c_ncu.log 2>&1
cut -c1-300 gpurun_out/r2c_bench.json; grep -v "^\[W\|^W1\|^$" gpurun_out/r2c_bench.err | tail -25; cut -c1-250 gpurun_out/r2c_mcast_probe.jsonl; tail -3 gpurun_out/r2c_ncu.log
