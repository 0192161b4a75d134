#!/bin/bash
# N=2 call: GPU test-suite with two GPUs, bench --gpus 2 (parity leg included), multicast shape probe (forced team of two).
set -u
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2b_topo.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2b_pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --steps 20 --warmup 5 --verbose > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err; echo "bench rc=$?" >> gpurun_out/r2b_bench.err
timeout 300 python tools/mcast_probe.py --layers 4 --size 1073741824 > gpurun_out/r2b_mcast_probe.jsonl 2> gpurun_out/r2b_mcast_probe.err; echo "probe rc=$?" >> gpurun_out/r2b_mcast_probe.err
tail -4 gpurun_out/r2b_pytest.log; cut -c1-400 gpurun_out/r2b_bench.json; tail -8 gpurun_out/r2b_bench.err; cat gpurun_out/r2b_mcast_probe.jsonl | cut -c1-330
