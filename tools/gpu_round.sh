#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; cut -c1-200 gpurun_out/bench_default.json
timeout 300 python tools/attn_bench.py 2>&1 | cut -c1-110 | head -3
timeout 300 python tools/gemv_bench.py --qt nf4 --l2pf 0 2>&1 | tail -1 | cut -c1-100
