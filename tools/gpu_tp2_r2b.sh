#!/bin/bash
set -u
mkdir -p gpurun_out/records
timeout 900 python -u -m pytest tests/test_gpu_model.py -q -m gpu -k "tp2 or sequential" -s --timeout 400 > gpurun_out/records/pytest_tp2_multi_gpu.log 2>&1; echo "pytest tp2 rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/records/pytest_tp2_multi_gpu.log | grep -E "^\[tp|passed|failed|Error" | tail -10
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/records/tp2_bench20.json 2> gpurun_out/records/tp2_bench20.err; echo "tp2 bench rc=$?"
tail -1 gpurun_out/records/tp2_bench20.json | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().split('\n')[-1]); print('tp2', round(d['value'],1), round(d['e2e']['value'],1))"
