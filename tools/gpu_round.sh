#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q -k "gemv or tiled or stream or program" > gpurun_out/pytest_gemv.log 2>&1; echo "pytest rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_gemv.log | tail -3
timeout 300 python tools/gemv_bench.py --qt int4 --l2pf 0 > gpurun_out/gemv_int4.log 2>&1; tail -1 gpurun_out/gemv_int4.log | cut -c1-120; timeout 300 python tools/gemv_bench.py --qt nf4 --l2pf 0 > gpurun_out/gemv_nf4.log 2>&1; tail -1 gpurun_out/gemv_nf4.log | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['per_layer_us'], [(r['shape'].split()[0], r['chain_us'], r['chain_GBs']) for r in d['rows']])"
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; cut -c1-200 gpurun_out/bench_default.json
