#!/bin/bash
# ncu evidence for profiles/: launch list of two decode steps + one --set full capture of the hot kernels.
set -u
mkdir -p gpurun_out
KRE='regex:qgemv|attn_|argmax|embedding|sample|rmsnorm|rope_kv|qgemm'
timeout 900 python bench.py --no-cpu-baseline --steps 4 --warmup 3 > gpurun_out/prof_bench.json 2> gpurun_out/prof_bench.err || { echo "bench failed"; tail -5 gpurun_out/prof_bench.err; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KRE" --launch-skip 2000 -c 400 --csv --log-file gpurun_out/launches_r1f.csv python bench.py --no-cpu-baseline --steps 4 --warmup 3 > gpurun_out/ncu_launch.log 2>&1; echo "ncu launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:qgemv_stream --launch-skip 1400 -c 5 -o gpurun_out/prof_qgemv_r1f -f python bench.py --no-cpu-baseline --steps 4 --warmup 3 > gpurun_out/ncu_full.log 2>&1; echo "ncu full gemv rc=$?"
timeout 900 python tools/prefill_bench.py --layers 2 --reps 2 > gpurun_out/prefill_small.log 2>&1 || { echo "prefill failed"; tail -5 gpurun_out/prefill_small.log; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:attn_prefill -c 1 --launch-skip 2 -o gpurun_out/prof_attn_prefill_r1f -f python tools/prefill_bench.py --layers 2 --reps 1 > gpurun_out/ncu_full2.log 2>&1; echo "ncu full attn_prefill rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:attn_decode -c 3 --launch-skip 8 -o gpurun_out/prof_attn_decode_r1f -f python tools/attn_bench.py --reps 1 > gpurun_out/ncu_full3.log 2>&1; echo "ncu full attn_decode rc=$?"
ls -la gpurun_out/*.ncu-rep
