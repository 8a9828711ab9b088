#!/bin/bash
# round-2 records after the batched kernel (csrc/qgemv_nb.cu) and the captured short-prompt prefill: bench lines, workload lines, ncu evidence
set -u
mkdir -p gpurun_out/final gpurun_out/records
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/final/bench_20.json 2> gpurun_out/final/bench_20.err; echo "bench 20 rc=$?"
timeout 600 python bench.py > gpurun_out/final/bench_256.json 2> gpurun_out/final/bench_256.err; echo "bench 256 rc=$?"
for WL in llama2-13b-int4-bs1-ctx2048 llama2-13b-int4-bs4-ctx2048 llama2-13b-int4-bs16-ctx2048; do
  timeout 300 python bench.py --no-cpu-baseline --steps 64 --warmup 8 --workload $WL > gpurun_out/records/$WL.json 2> gpurun_out/records/$WL.err; echo "$WL rc=$?"
done
# launch list of the batch-16 step (after the plain run above exited 0): which kernels a step is made of
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/records/launches_13b_bs16.csv \
  python bench.py --no-cpu-baseline --steps 2 --warmup 3 --workload llama2-13b-int4-bs16-ctx2048 > gpurun_out/records/ncu_launches.log 2>&1; echo "ncu launch list rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:qgemv_nb --launch-skip 3 -c 1 -o gpurun_out/records/qgemv_nb_full -f \
  python tools/gemv_bench.py --model Llama-2-13b-hf --qt int4 --batch 16 --json --reps 2 --only fc_1 > gpurun_out/records/ncu_full.log 2>&1; echo "ncu full rc=$?"
python - <<'P'
import json
for f in ["final/bench_20","final/bench_256","records/llama2-13b-int4-bs1-ctx2048","records/llama2-13b-int4-bs4-ctx2048","records/llama2-13b-int4-bs16-ctx2048"]:
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().split("\n")[-1])
        r=d.get("roofline") or {}
        print(f, round(d["value"],2), "tok/s", round(d["ms_per_step"],3), "ms", "e2e", (d.get("e2e") or {}).get("value"), "frac", r.get("frac"), "step", r.get("whole_step_frac"), "clocks", (d.get("clocks") or {}).get("sm_mhz"), (d.get("clocks") or {}).get("samples"), "cpu", (d.get("cpu_baseline") or {}).get("value"))
    except Exception as e: print(f, "ERR", e)
P
