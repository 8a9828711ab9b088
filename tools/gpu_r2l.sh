#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 300 python -u -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "tiled or qgemv" --timeout 40 > gpurun_out/pytest_k.log 2>&1; echo "pytest rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_k.log | grep -E "passed|failed|Error|assert|Timeout" | tail -5
for b in 4 8 16; do
  timeout 120 python tools/gemv_bench.py --model Llama-2-13b-hf --qt int4 --batch $b --json --reps 10 > gpurun_out/gemv13_b$b.json 2> gpurun_out/gemv13_b$b.err; echo "b=$b rc=$?"
done
python - <<'P'
import json
for f in ["gemv13_b4","gemv13_b8","gemv13_b16"]:
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().split("\n")[-1])
        print(f, d["per_layer_us"], [(r["shape"][:10], r["chain_us"], r["chain_GBs"]) for r in d["rows"]])
    except Exception as e: print(f, "ERR", e)
P
