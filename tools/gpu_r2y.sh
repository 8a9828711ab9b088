#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -u -m pytest tests/test_gpu_model.py tests/test_gpu_sample.py -x -q -m gpu --timeout 120 > gpurun_out/pytest_m.log 2>&1; echo "pytest rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_m.log | grep -E "passed|failed|Error|assert|Timeout" | tail -12
timeout 300 python bench.py --no-cpu-baseline --steps 20 --warmup 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().split('\n')[-1]); print('bench20', round(d['value'],1), round(d['e2e']['value'],1))"
