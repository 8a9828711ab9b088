#!/bin/bash
timeout 300 python -u -m pytest tests/test_gpu_kernels.py -x -q -m gpu --timeout 60 -k "attn" 2>&1 | grep -E "passed|failed" | tail -2
timeout 200 python -u -m pytest tests/test_gpu_model.py -x -q -m gpu --timeout 120 -k "13b or batched" 2>&1 | grep -E "passed|failed" | tail -2
timeout 300 python bench.py --no-cpu-baseline --steps 32 --warmup 4 --workload llama2-13b-int4-bs16-ctx2048 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().split('\n')[-1]); print('13b bs16', round(d['value'],1), d['ms_per_step'], d['roofline']['whole_step_frac'])"
