#!/bin/bash
# round-2 final validation: what the driver runs at round end (GPU tests, smoke, the bench line) + a refreshed launch list of the headline step
set -u
mkdir -p gpurun_out/final gpurun_out/records
timeout 1200 python -m pytest tests/ -x -q -m gpu --timeout 300 > gpurun_out/final/pytest_gpu.log 2>&1; echo "pytest -m gpu rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/final/pytest_gpu.log | tail -3
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('__SMOKE_OK__')" > gpurun_out/final/smoke.log 2>&1; echo "smoke rc=$?"; grep -v "^DEBUG" gpurun_out/final/smoke.log | tail -2
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/final/bench_20.json 2> gpurun_out/final/bench_20.err; echo "bench 20 rc=$?"
KRE='regex:qgemv|attn_|argmax|embedding|sample|rmsnorm|rope_kv|qgemm'
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KRE" --launch-skip 2000 -c 326 --csv --log-file gpurun_out/records/launches_r02.csv python bench.py --no-cpu-baseline --steps 4 --warmup 3 > gpurun_out/records/ncu_launch.log 2>&1; echo "ncu launch list rc=$?"
python - <<'P'
import json
for f in ["final/bench_20"]:
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().split("\n")[-1])
        r=d.get("roofline") or {}
        print(f, round(d["value"],2), "tok/s", round(d["ms_per_step"],3), "ms", "e2e", (d.get("e2e") or {}).get("value"), "frac", r.get("frac"), "step", r.get("whole_step_frac"), "clocks", (d.get("clocks") or {}).get("sm_mhz"), (d.get("clocks") or {}).get("samples"), "cpu", (d.get("cpu_baseline") or {}).get("value"))
    except Exception as e: print(f, "ERR", e)
P
