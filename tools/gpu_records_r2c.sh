#!/bin/bash
set -u
mkdir -p gpurun_out/records
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -k 'regex:qgemv|attn_decode|rmsnorm|argmax|embedding_kernel' --launch-skip 400 -c 900 --csv --log-file gpurun_out/records/launches_13b_bs16.csv \
  python bench.py --no-cpu-baseline --steps 2 --warmup 3 --workload llama2-13b-int4-bs16-ctx2048 > gpurun_out/records/ncu_launches.log 2>&1; echo "ncu launch list rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:qgemv_nb --launch-skip 3 -c 1 -o gpurun_out/records/qgemv_nb_full -f \
  python tools/gemv_bench.py --model Llama-2-13b-hf --qt int4 --batch 16 --json --reps 2 --only fc_1 > gpurun_out/records/ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out/records/*.ncu-rep
timeout 120 python tools/gemv_bench.py --model Llama-2-13b-hf --qt int4 --batch 16 --json --reps 10 > gpurun_out/records/gemv13_b16.json 2>/dev/null
timeout 120 python tools/gemv_bench.py --model Llama-2-7b-hf --qt nf4 --batch 16 --json --reps 10 > gpurun_out/records/gemv7_b16.json 2>/dev/null
python - <<'P'
import json
for f in ["gemv7_b16","gemv13_b16"]:
    try:
        d=json.loads(open(f"gpurun_out/records/{f}.json").read().strip().split("\n")[-1])
        print(f, d["per_layer_us"], [(r["shape"][:10], r["chain_us"], r["chain_GBs"]) for r in d["rows"]])
    except Exception as e: print(f, "ERR", e)
P
