#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_20.json 2> gpurun_out/bench_20.err; echo "bench rc=$?"; python - <<'P'
import json
d=json.loads(open("gpurun_out/bench_20.json").read().strip().split("\n")[-1])
print({k:d[k] for k in ("value","ms_per_step","clocks","e2e","gpu_launches")}); print(d["roofline"])
P
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_256.json 2> gpurun_out/bench_256.err; echo "bench rc=$?"; python - <<'P'
import json
d=json.loads(open("gpurun_out/bench_256.json").read().strip().split("\n")[-1])
print({k:d[k] for k in ("value","ms_per_step","clocks","e2e","gpu_launches")}); print(d["roofline"])
P
timeout 600 python -m pytest tests/test_gpu_model.py -x -q -m gpu -k "llama7b" > gpurun_out/pytest_7b.log 2>&1; echo "7b rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_7b.log | tail -3
