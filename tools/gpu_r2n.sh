#!/bin/bash
set -u
mkdir -p gpurun_out
for occ in 0 14 0 14; do
B200_TUNE_OCC=$occ timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_occ$occ.json 2> gpurun_out/bench_occ$occ.err; python - <<P
import json
d=json.loads(open("gpurun_out/bench_occ$occ.json").read().strip().split("\n")[-1])
print("occ $occ", round(d["value"],1), round(d["ms_per_step"],4), round(d["e2e"]["value"],1))
P
done
B200_TUNE_OCC=0 timeout 300 python bench.py --no-cpu-baseline --workload llama2-7b-nf4-bs1-ctx2048 --steps 128 > gpurun_out/b2k_0.json 2>/dev/null; B200_TUNE_OCC=14 timeout 300 python bench.py --no-cpu-baseline --workload llama2-7b-nf4-bs1-ctx2048 --steps 128 > gpurun_out/b2k_14.json 2>/dev/null
python - <<P
import json
for f in ("b2k_0","b2k_14"):
    d=json.loads(open(f"gpurun_out/{f}.json").read().strip().split("\n")[-1]); print(f, round(d["value"],1), round(d["ms_per_step"],4))
P
timeout 300 python -u -m pytest tests/test_gpu_model.py -x -q -m gpu --timeout 120 -k "toy or llama7b" > gpurun_out/pytest_m.log 2>&1; echo "pytest rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_m.log | grep -E "passed|failed|Error" | tail -3
