#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_model.py -x -q -m gpu > gpurun_out/pytest_kern.log 2>&1; echo "pytest rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_kern.log | tail -4
timeout 600 python tools/trace_step.py --layers 4 --prologue > gpurun_out/trace_prologue_nf4.txt 2>&1; echo "trace rc=$?"; tail -18 gpurun_out/trace_prologue_nf4.txt
timeout 600 python tools/trace_step.py --layers 4 > gpurun_out/trace_nf4.txt 2>&1; echo "trace rc=$?"; tail -18 gpurun_out/trace_nf4.txt
timeout 300 python tools/gemv_bench.py --qt nf4 --json > gpurun_out/gemv_v4_nf4.json 2> gpurun_out/gemv_v4_nf4.err
timeout 300 python tools/gemv_bench.py --qt int4 --json > gpurun_out/gemv_v4_int4.json 2> gpurun_out/gemv_v4_int4.err
python - <<'P'
import json
for f in ["gemv_v4_nf4","gemv_v4_int4"]:
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().split("\n")[-1])
        print(f, d["per_layer_us"], d["step_est_us"], [(r["shape"][:10], r["chain_us"], r["chain_GBs"], r["cold_us"]) for r in d["rows"]])
    except Exception as e: print(f, "ERR", e)
P
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/bench_default.json
