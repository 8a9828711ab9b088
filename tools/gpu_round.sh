#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_decode_cluster_kernel -c 2 --launch-skip 4 -o gpurun_out/prof_attn_cluster_r1g -f python tools/attn_bench.py --reps 1 > gpurun_out/ncu_a1.log 2>&1; echo "rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_decode_gqa_kernel -c 2 --launch-skip 4 -o gpurun_out/prof_attn_gqa_r1g -f python tools/attn_bench.py --reps 1 > gpurun_out/ncu_a2.log 2>&1; echo "rc=$?"
