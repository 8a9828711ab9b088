#!/bin/bash
timeout 300 python -u -m pytest tests/test_gpu_kernels.py -x -q -m gpu --timeout 60 -k "qgemm" 2>&1 | grep -E "passed|failed|Error|assert" | tail -5
timeout 200 python tools/gemm_bench.py 2>&1 | grep -v "^DEBUG\|Warn\|warn" | tail -4 | cut -c1-160
timeout 200 python tools/gemm_bench.py --model Mistral-7B-v0.1 --qt int4 2>&1 | grep -v "^DEBUG\|Warn\|warn" | tail -4 | cut -c1-160
timeout 300 python tools/prefill_bench.py 2>&1 | grep -v "^DEBUG\|Warn\|warn" | tail -1
timeout 300 python tools/prefill_bench.py --model Llama-2-7b-hf --qt nf4 2>&1 | grep -v "^DEBUG\|Warn\|warn" | tail -1
