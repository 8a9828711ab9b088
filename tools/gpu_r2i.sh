#!/bin/bash
set -u
mkdir -p gpurun_out/records
timeout 900 python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "tiled" > gpurun_out/pytest_tiled.log 2>&1; echo "tiled rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_tiled.log | tail -12
for WL in llama2-13b-int4-bs16-ctx2048 llama2-13b-int4-bs4-ctx2048; do
  timeout 600 python bench.py --no-cpu-baseline --steps 64 --warmup 8 --workload $WL > gpurun_out/records/$WL.json 2> gpurun_out/records/$WL.err; echo "$WL rc=$?"; tail -3 gpurun_out/records/$WL.err
done
python - <<'P'
import json,glob
for f in sorted(glob.glob("gpurun_out/records/llama2-13b*.json")):
    try:
        d=json.loads(open(f).read().strip().split("\n")[-1])
        print(f.split("/")[-1], round(d["value"],1), "tok/s", round(d["ms_per_step"],3), "ms", "gemv frac", round(d["roofline"]["frac"],3), "step frac", round(d["roofline"]["whole_step_frac"],3))
    except Exception as e: print(f, "ERR", e)
P
