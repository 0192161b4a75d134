#!/bin/bash
# N=2 call: push/pull mix on one port; ncu (application replay) of a push, a pull and a relay launch; config_bench on ONE GPU
# (replaces the stale round-1 file).
set -u
mkdir -p gpurun_out
timeout 300 python tools/p2p_mix_probe.py --layers 8 --size 1073741824 > gpurun_out/r2f_mix.jsonl 2> gpurun_out/r2f_mix.err; echo "rc=$?" >> gpurun_out/r2f_mix.err
CUDA_VISIBLE_DEVICES=0 timeout 400 python tools/config_bench.py --configs readme,ring64,seed64,llama8b --reps 5 > gpurun_out/r2f_config_n1.jsonl 2> gpurun_out/r2f_config_n1.err; echo "rc=$?" >> gpurun_out/r2f_config_n1.err
SECS="--section SpeedOfLight --section WarpStateStats --section SchedulerStats --section MemoryWorkloadAnalysis --section LaunchStats --section Occupancy"
for CASE in push pull relay; do
  ONE="python tools/p2p_mix_probe.py --layers 2 --size 536870912 --one $CASE"
  timeout 120 $ONE > gpurun_out/r2f_one_$CASE.log 2>&1 &&
  timeout 600 ncu --replay-mode application $SECS --clock-control none -k regex:copy_csum -c 12 --csv --page raw --log-file gpurun_out/r2f_ncu_$CASE.csv $ONE > gpurun_out/r2f_ncu_$CASE.log 2>&1
  echo "ncu $CASE rc=$?" >> gpurun_out/r2f_ncu_$CASE.log
done
cut -c1-200 gpurun_out/r2f_mix.jsonl; cut -c1-260 gpurun_out/r2f_config_n1.jsonl | head -20; tail -2 gpurun_out/r2f_ncu_*.log
