#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2m_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2m_pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29601 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2m_bench.json 2> gpurun_out/r2m_bench.err; echo "bench rc=$?" >> gpurun_out/r2m_bench.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29602 bench.py --impl reference --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2m_ref.json 2> gpurun_out/r2m_ref.err; echo "ref rc=$?" >> gpurun_out/r2m_ref.err
tail -n 3 gpurun_out/r2m_pytest.log; grep '^{' gpurun_out/r2m_bench.json | cut -c1-200; tail -n 1 gpurun_out/r2m_bench.err; grep '^{' gpurun_out/r2m_ref.json | cut -c1-200; tail -n 1 gpurun_out/r2m_ref.err
