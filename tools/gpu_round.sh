#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
for WL in mistral-7b-int4-bs1 llama2-7b-nf4-bs1; do
timeout 600 python bench.py --no-cpu-baseline --steps 128 --workload $WL > gpurun_out/bench_$WL.json 2> gpurun_out/bench_$WL.err; echo "bench $WL rc=$?"; cut -c1-200 gpurun_out/bench_$WL.json
done
