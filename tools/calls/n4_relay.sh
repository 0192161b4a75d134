#!/bin/bash
# N=4: relay at SMALL layers (64 MiB): one process with event-ordered rounds vs 4 processes with a barrier per round vs 4 processes
# with device-side waits.  Also at 436 MB layers.
set -u
mkdir -p gpurun_out
for SZ in 67108864 436224000; do
timeout 200 python tools/fanout_bench.py --one-seed --layers 32 --size $SZ --reps 5 --cases relay4p,relay8p > gpurun_out/r2j_sp_$SZ.jsonl 2> gpurun_out/r2j_sp_$SZ.err
for P in 4 8; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29591 tools/fanout_bench_mp.py --one-seed --layers 32 --size $SZ --pieces $P --reps 5 > gpurun_out/r2j_mp_${SZ}_$P.jsonl 2> gpurun_out/r2j_mp_${SZ}_$P.err
done
done
cat gpurun_out/r2j_*.jsonl | cut -c1-420; tail -n 3 gpurun_out/r2j_*.err | grep -v "^\[W\|^W1\|OMP\|\*\*\*\|^$" | head -20
