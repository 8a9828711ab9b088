#!/bin/bash
# single-GPU records of BASELINE.json configs 2-4 (one JSON line per workload under gpurun_out/records/)
set -u
mkdir -p gpurun_out/records
for WL in llama2-7b-nf4-bs1-ctx2048 llama2-13b-int4-bs1-ctx2048 llama2-13b-int4-bs4-ctx2048 llama2-13b-int4-bs16-ctx2048 mistral-7b-int4-bs1 mistral-7b-int4-bs1-ctx3936; do
  timeout 600 python bench.py --no-cpu-baseline --steps 128 --warmup 8 --workload $WL > gpurun_out/records/$WL.json 2> gpurun_out/records/$WL.err; echo "$WL rc=$?"
done
timeout 600 python tools/prefill_bench.py --model Mistral-7B-v0.1 --qt int4 --tokens 4096 > gpurun_out/records/prefill_mistral-7b-int4-4096.json 2> gpurun_out/records/prefill_mistral.err; echo "prefill mistral rc=$?"
timeout 600 python tools/prefill_bench.py --model Llama-2-7b-hf --qt nf4 --tokens 4096 > gpurun_out/records/prefill_llama2-7b-nf4-4096.json 2> gpurun_out/records/prefill_llama.err; echo "prefill llama rc=$?"
python - <<'P'
import json,glob
for f in sorted(glob.glob("gpurun_out/records/*.json")):
    try:
        d=json.loads(open(f).read().strip().split("\n")[-1])
        if "roofline" in d: print(f.split("/")[-1], round(d["value"],1), "tok/s", round(d["ms_per_step"],3), "ms", "gemv frac", round(d["roofline"]["frac"],3), "step frac", round(d["roofline"]["whole_step_frac"],3), d["clocks"].get("sm_mhz"))
        else: print(f.split("/")[-1], {k:(round(v,2) if isinstance(v,float) else v) for k,v in d.items() if not isinstance(v,(dict,list))})
    except Exception as e: print(f, "ERR", e)
P
