#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python tools/trace_step.py --layers 4 --prologue > gpurun_out/trace_prologue_nf4.txt 2>&1; echo "trace rc=$?"; tail -20 gpurun_out/trace_prologue_nf4.txt
for occ in 10 11; do
  timeout 300 python tools/gemv_bench.py --qt nf4 --json --occ $occ > gpurun_out/gemv_v4_occ$occ.json 2> gpurun_out/gemv_v4_occ$occ.err; echo "occ $occ rc=$?"
done
python - <<'P'
import json
for f in ["gemv_v4_occ10","gemv_v4_occ11"]:
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().split("\n")[-1])
        print(f, d["per_layer_us"], d["step_est_us"], [(r["shape"][:10], r["chain_us"], r["chain_GBs"], r["cold_us"]) for r in d["rows"]])
    except Exception as e: print(f, "ERR", e)
P
