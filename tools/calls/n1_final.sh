#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r2h_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2h_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2h_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r2h_smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2h_bench.json 2> gpurun_out/r2h_bench.err; echo "bench rc=$?" >> gpurun_out/r2h_bench.err
PROF="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 300 $PROF > gpurun_out/r2h_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r2h_launches.csv $PROF > gpurun_out/r2h_ncu1.log 2>&1
tail -n 3 gpurun_out/r2h_pytest.log; cat gpurun_out/r2h_smoke.log; cut -c1-300 gpurun_out/r2h_bench.json; tail -n 2 gpurun_out/r2h_bench.err
