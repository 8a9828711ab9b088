#!/bin/bash
# round-2 ncu evidence for profiles/: launch list of decode steps + one --set full capture of the dominant kernel
set -u
mkdir -p gpurun_out
KRE='regex:qgemv|attn_|argmax|embedding|sample|rmsnorm|rope_kv|qgemm'
timeout 600 python bench.py --no-cpu-baseline --steps 4 --warmup 3 > gpurun_out/prof_bench.json 2> gpurun_out/prof_bench.err || { echo "bench failed"; tail -5 gpurun_out/prof_bench.err; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KRE" --launch-skip 2000 -c 400 --csv --log-file gpurun_out/launches_r02.csv python bench.py --no-cpu-baseline --steps 4 --warmup 3 > gpurun_out/ncu_launch.log 2>&1; echo "ncu launch list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:qgemv_v4 --launch-skip 1400 -c 5 -o gpurun_out/prof_qgemv_r02 -f python bench.py --no-cpu-baseline --steps 4 --warmup 3 > gpurun_out/ncu_full.log 2>&1; echo "ncu full gemv rc=$?"
ls -la gpurun_out/*.ncu-rep gpurun_out/launches_r02.csv
