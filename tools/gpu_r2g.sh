#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_gpu.log | tail -5
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_20.json 2> gpurun_out/bench_20.err; echo "bench rc=$?"; python - <<'P'
import json
d=json.loads(open("gpurun_out/bench_20.json").read().strip().split("\n")[-1])
print({k:d[k] for k in ("value","ms_per_step","clocks","e2e","gpu_launches")}); print(d["roofline"]); print(d["cpu_baseline"])
P
timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cut -c1-1200 gpurun_out/bench_ref.json
nproc; free -g | head -2
