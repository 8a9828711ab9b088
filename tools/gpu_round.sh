#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:qgemv_stream --launch-skip 3 -c 1 -o gpurun_out/prof_qgemv_int4 -f python tools/gemv_bench.py --qt int4 --l2pf 0 --reps 2 > gpurun_out/ncu_int4.log 2>&1; echo "rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:qgemv_stream --launch-skip 3 -c 1 -o gpurun_out/prof_qgemv_nf4 -f python tools/gemv_bench.py --qt nf4 --l2pf 0 --reps 2 > gpurun_out/ncu_nf4.log 2>&1; echo "rc=$?"
