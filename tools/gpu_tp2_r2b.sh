#!/bin/bash
set -u
mkdir -p gpurun_out/records
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/records/tp2_bench20.json 2> gpurun_out/records/tp2_bench20.err; echo "tp2 bench rc=$?"
tail -1 gpurun_out/records/tp2_bench20.json | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().split('\n')[-1]); print('tp2', round(d['value'],1), round(d['e2e']['value'],1), d['config'].get('parallelism'), d['clocks'].get('samples'))"
