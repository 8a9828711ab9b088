#!/bin/bash
set -u
mkdir -p gpurun_out
KRE='regex:qgemv|attn_|argmax|embedding|sample|rmsnorm|rope_kv|qgemm'
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KRE" --csv --log-file gpurun_out/launches_r02_all.csv python bench.py --no-cpu-baseline --steps 4 --warmup 3 > gpurun_out/ncu_launch.log 2>&1; echo "ncu launch list rc=$?"
wc -l gpurun_out/launches_r02_all.csv
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_decode --launch-skip 40 -c 2 -o gpurun_out/prof_attn_r02 -f python bench.py --no-cpu-baseline --steps 4 --warmup 3 > gpurun_out/ncu_full_attn.log 2>&1; echo "ncu full attn rc=$?"
