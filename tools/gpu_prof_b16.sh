#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 300 ncu --set full --clock-control none --import-source on -k regex:qgemv_nb --launch-skip 3 -c 1 -o gpurun_out/prof_nb -f python tools/gemv_bench.py --model Llama-2-13b-hf --qt int4 --batch 16 --json --reps 2 --only mlp.proj > gpurun_out/ncu_nb.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/prof_nb.ncu-rep
