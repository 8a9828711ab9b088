#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 400 python -u -m pytest tests/test_gpu_kernels.py -x -q -m gpu --timeout 40 -k "tiled or qgemv" > gpurun_out/pytest_k.log 2>&1; echo "kernels rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_k.log | grep -E "passed|failed|Error|assert|Timeout" | tail -8
timeout 120 python tools/gemv_bench.py --model Llama-2-7b-hf --qt nf4 --batch 16 --json --reps 10 > gpurun_out/gemv7_b16.json 2>/dev/null
timeout 120 python tools/gemv_bench.py --model Llama-2-13b-hf --qt int4 --batch 16 --json --reps 10 > gpurun_out/gemv13_b16.json 2>/dev/null
python - <<'P'
import json
for f in ["gemv7_b16","gemv13_b16"]:
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().split("\n")[-1])
        print(f, d["per_layer_us"], [(r["shape"][:10], r["chain_us"], r["chain_GBs"]) for r in d["rows"]])
    except Exception as e: print(f, "ERR", e)
P
