#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q -k "int4 or program or tp2 or model" > gpurun_out/pytest_int4.log 2>&1; echo "pytest rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_int4.log | tail -4
timeout 300 python tools/gemv_bench.py --qt int4 --l2pf 0 > gpurun_out/gemv_int4.log 2>&1; tail -1 gpurun_out/gemv_int4.log | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['per_layer_us'], [(r['shape'].split()[0], r['chain_us'], r['chain_GBs']) for r in d['rows']])"
for WL in llama2-13b-int4-bs1-ctx2048 mistral-7b-int4-bs1; do
timeout 600 python bench.py --no-cpu-baseline --steps 128 --workload $WL > gpurun_out/bench_$WL.json 2> gpurun_out/bench_$WL.err; cut -c1-150 gpurun_out/bench_$WL.json
done
