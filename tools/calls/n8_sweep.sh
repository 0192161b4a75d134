#!/bin/bash
set -u
mkdir -p gpurun_out
i=0
for OPT in "--ctas-per-sm 1" "--tma-cfg 1" "--tma-cfg 2" "--variant 0" ""; do
i=$((i+1))
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29630+i)) bench.py --gpus 8 --steps 50 --warmup 5 $OPT --no-north-star --no-parity --no-e2e --no-extras > gpurun_out/r2q_$i.json 2> gpurun_out/r2q_$i.err
echo "$OPT: $(grep '^{' gpurun_out/r2q_$i.json | python -c "import sys,json; j=json.loads(sys.stdin.read()); print(j['ms_per_step'], j['device_ms_per_step'], j['completion_device_ms'], j['roofline']['frac'])")"
done
