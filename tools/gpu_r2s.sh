#!/bin/bash
set -u
mkdir -p gpurun_out/records
timeout 400 python -u -m pytest tests/test_gpu_model.py tests/test_gpu_checkpoint.py -x -q -m gpu --timeout 180 > gpurun_out/pytest_m.log 2>&1; echo "model rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_m.log | grep -E "passed|failed|Error|assert" | tail -4
timeout 120 python tools/prefill16_profile.py 2>&1 | grep -v "^DEBUG" | grep -E "prefill\(16\)|qgemv|swiglu|rmsnorm|aten::add|index_select" | cut -c1-200
for m in "Mistral-7B-v0.1 int4" "Llama-2-7b-hf nf4"; do set -- $m; timeout 300 python tools/prefill_bench.py --model $1 --qt $2 --tokens 4096 2>/dev/null | tail -1 | cut -c1-300; done
timeout 300 python bench.py --no-cpu-baseline --steps 20 --warmup 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().split('\n')[-1]); print('bench20', round(d['value'],1), round(d['e2e']['value'],1))"
