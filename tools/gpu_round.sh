#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q -k "attn or program or model or sample" > gpurun_out/pytest_attn.log 2>&1; echo "pytest rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_attn.log | tail -3
timeout 300 python tools/attn_bench.py 2>&1 | cut -c1-110 | head -3
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; cut -c1-200 gpurun_out/bench_default.json
