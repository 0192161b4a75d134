#!/bin/bash
# N=2 call: GPU suite again (store-mode sums, fixed ordering test), bench, NVLink counters per way of moving bytes, ncu on a multicast
# launch in application-replay mode (kernel replay cannot save/restore multicast mappings).
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2e_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2e_pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 2 --steps 20 --warmup 5 --verbose > gpurun_out/r2e_bench.json 2> gpurun_out/r2e_bench.err; echo "bench rc=$?" >> gpurun_out/r2e_bench.err
timeout 300 python tools/nvlink_counters.py > gpurun_out/r2e_nvlink_counters.jsonl 2> gpurun_out/r2e_nvlink_counters.err; echo "rc=$?" >> gpurun_out/r2e_nvlink_counters.err
ONE="python tools/mcast_probe.py --layers 2 --size 536870912 --one"
timeout 120 $ONE > gpurun_out/r2e_one_plain.log 2>&1 &&
timeout 900 ncu --replay-mode application --section SpeedOfLight --section WarpStateStats --section SchedulerStats --section MemoryWorkloadAnalysis --section LaunchStats --section Occupancy --clock-control none -k regex:copy_csum_kernel -c 1 --csv --page raw --log-file gpurun_out/r2e_mcast_ncu_raw.csv $ONE > gpurun_out/r2e_ncu.log 2>&1
echo "ncu rc=$?" >> gpurun_out/r2e_ncu.log
tail -3 gpurun_out/r2e_pytest.log; cut -c1-300 gpurun_out/r2e_bench.json; grep -v "^\[W\|^W1\|^$" gpurun_out/r2e_bench.err | tail -8; cut -c1-600 gpurun_out/r2e_nvlink_counters.jsonl; tail -5 gpurun_out/r2e_ncu.log; head -c 600 gpurun_out/r2e_mcast_ncu_raw.csv
