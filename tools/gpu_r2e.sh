#!/bin/bash
set -u
mkdir -p gpurun_out
for occ in 0 13; do
  timeout 300 python tools/gemv_bench.py --qt int4 --json --occ $occ > gpurun_out/gemv_int4_occ$occ.json 2> gpurun_out/gemv_int4_occ$occ.err; echo "occ $occ rc=$?"; tail -2 gpurun_out/gemv_int4_occ$occ.err
done
python - <<'P'
import json
for f in ["gemv_int4_occ0","gemv_int4_occ13"]:
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().split("\n")[-1])
        print(f, d["per_layer_us"], d["step_est_us"], [(r["shape"][:10], r["chain_us"], r["chain_GBs"], r["cold_us"]) for r in d["rows"]])
    except Exception as e: print(f, "ERR", e)
P
