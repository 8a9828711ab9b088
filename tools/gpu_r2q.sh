#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 200 python tools/trace_step.py --layers 4 > gpurun_out/trace_fa.txt 2>&1; echo "trace rc=$?"; tail -19 gpurun_out/trace_fa.txt | cut -c1-260
B200_FUSE_ATTN_MAX_SEQ=1024 timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_fa.json 2> gpurun_out/bench_fa.err; python - <<P
import json
d=json.loads(open("gpurun_out/bench_fa.json").read().strip().split("\n")[-1])
print("fused", round(d["value"],1), round(d["ms_per_step"],4), round(d["e2e"]["value"],1), d["launches_per_step"])
P
