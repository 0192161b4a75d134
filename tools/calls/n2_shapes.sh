#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 400 python tools/push_shape_probe.py > gpurun_out/r2o_push_shapes.jsonl 2> gpurun_out/r2o_push_shapes.err; echo "rc=$?" >> gpurun_out/r2o_push_shapes.err
python -c "
import json
for l in open('gpurun_out/r2o_push_shapes.jsonl'):
    if l.startswith('{'):
        j=json.loads(l); print(j.get('case'), j.get('ring'), j.get('ctas_per_sm'), round(j.get('GBps_per_direction',0),1), j.get('self_check'))
"; tail -n 2 gpurun_out/r2o_push_shapes.err
