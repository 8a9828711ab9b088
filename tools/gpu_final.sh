#!/bin/bash
# Round-end evidence on one GPU: smoke, all GPU tests, the bench line, the reference arm, the other workloads, ncu.
set -u
mkdir -p gpurun_out
timeout 600 python __graft_entry__.py > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_gpu.log | tail -2
timeout 900 python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench rc=$?"; cat gpurun_out/bench_full.json
timeout 900 python bench.py --impl reference --steps 8 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cut -c1-200 gpurun_out/bench_ref.json
for WL in llama2-7b-nf4-bs1-ctx2048 llama2-13b-int4-bs1-ctx2048 llama2-13b-int4-bs4-ctx2048 llama2-13b-int4-bs16-ctx2048 mistral-7b-int4-bs1 mistral-7b-int4-bs1-ctx3936; do
timeout 600 python bench.py --no-cpu-baseline --steps 128 --workload $WL > gpurun_out/bench_$WL.json 2> gpurun_out/bench_$WL.err; echo "$WL rc=$?"; cut -c1-140 gpurun_out/bench_$WL.json
done
timeout 600 python tools/prefill_bench.py > gpurun_out/prefill_mistral.log 2>&1; tail -1 gpurun_out/prefill_mistral.log
timeout 600 python tools/prefill_bench.py --model Llama-2-7b-hf --qt nf4 > gpurun_out/prefill_llama.log 2>&1; tail -1 gpurun_out/prefill_llama.log
timeout 300 python tools/gemv_bench.py --qt nf4 --l2pf 0 > gpurun_out/gemv_nf4.log 2>&1
timeout 300 python tools/gemv_bench.py --qt int4 --l2pf 0 > gpurun_out/gemv_int4.log 2>&1
timeout 300 python tools/attn_bench.py > gpurun_out/attn_bench.log 2>&1
KRE='regex:qgemv|attn_|argmax|embedding|sample|rmsnorm|rope_kv|qgemm'
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KRE" --launch-skip 600 -c 400 --csv --log-file gpurun_out/launches_r1g.csv python bench.py --no-cpu-baseline --steps 4 --warmup 3 > gpurun_out/ncu_launch.log 2>&1; echo "ncu launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:qgemv_stream --launch-skip 1400 -c 5 -o gpurun_out/prof_qgemv_r1g -f python bench.py --no-cpu-baseline --steps 4 --warmup 3 > gpurun_out/ncu_full.log 2>&1; echo "ncu full gemv rc=$?"
