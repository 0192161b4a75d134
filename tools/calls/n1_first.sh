#!/bin/bash
# N=1 call: GPU test-suite, default bench line, reference arm, ncu launch list + full capture of the copy kernel (config 3).
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/r2a_gpus.txt 2>&1
nproc >> gpurun_out/r2a_gpus.txt; free -g >> gpurun_out/r2a_gpus.txt
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2a_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 --verbose > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench rc=$?" >> gpurun_out/r2a_bench.err
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2a_ref.json 2> gpurun_out/r2a_ref.err; echo "ref rc=$?" >> gpurun_out/r2a_ref.err
PROF="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 300 $PROF > gpurun_out/r2a_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r2a_launches.csv $PROF > gpurun_out/r2a_ncu1.log 2>&1
timeout 300 $PROF > gpurun_out/r2a_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:copy_csum_tma -s 2 -c 2 -o gpurun_out/r2a_prof $PROF > gpurun_out/r2a_ncu2.log 2>&1
tail -3 gpurun_out/r2a_pytest.log; cat gpurun_out/r2a_bench.json | cut -c1-600; tail -5 gpurun_out/r2a_bench.err
