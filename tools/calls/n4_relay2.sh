#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "relay or wait or relaunch" > gpurun_out/r2k_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2k_pytest.log
for SZ in 67108864 436224000; do
timeout 200 python tools/fanout_bench.py --one-seed --layers 32 --size $SZ --reps 5 --cases relay4p,relay4pf,relay8p,relay8pf,relay16pf > gpurun_out/r2k_sp_$SZ.jsonl 2> gpurun_out/r2k_sp_$SZ.err
done
tail -n 3 gpurun_out/r2k_pytest.log; cat gpurun_out/r2k_sp_*.jsonl | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        j=json.loads(l); print(j['layer_bytes'], j['case'], round(j['completion_ms_best'],3), j['verified'])
"; tail -n 3 gpurun_out/r2k_sp_*.err
