#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu -k "multicast or mcast or wait or relay or vmm" > gpurun_out/r2l_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2l_pytest.log
timeout 300 python tools/mcast_probe.py --layers 4 --size 1073741824 --quick > gpurun_out/r2l_mcast_probe.jsonl 2> gpurun_out/r2l_mcast_probe.err; echo "rc=$?" >> gpurun_out/r2l_mcast_probe.err
tail -n 3 gpurun_out/r2l_pytest.log; python -c "
import sys,json
for l in open('gpurun_out/r2l_mcast_probe.jsonl'):
    if l.startswith('{'):
        j=json.loads(l); print(j['case'], j['variant'], j['reg_cfg'], j['tma_cfg'], round(j['seed_egress_GBps'],1), j['verified'])
"
