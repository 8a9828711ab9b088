#!/bin/bash
# round-2 first measurement: probe, GPU tests, stream-K GEMV (v4) against the round-1 kernel, the bench line, launch floor
set -u
mkdir -p gpurun_out
(python -c "import bitsandbytes" 2>&1 | tail -1; pip list 2>/dev/null | grep -i -E "bitsandbytes|auto.gptq|torchao"; find / -iname "*bitsandbytes*.whl" 2>/dev/null | head -3; date -u) > gpurun_out/bnb_probe.txt 2>&1
cat gpurun_out/bnb_probe.txt
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_gpu.log | tail -6
for qt in nf4 int4; do
  timeout 300 python tools/gemv_bench.py --qt $qt --json --tiny > gpurun_out/gemv_v4_$qt.json 2> gpurun_out/gemv_v4_$qt.err; echo "v4 $qt rc=$?"
  timeout 300 python tools/gemv_bench.py --qt $qt --json --tiny --depth 98 > gpurun_out/gemv_old_$qt.json 2> gpurun_out/gemv_old_$qt.err; echo "old $qt rc=$?"
  timeout 300 python tools/gemv_bench.py --qt $qt --json --tiny --depth 99 > gpurun_out/gemv_v4dbg_$qt.json 2> gpurun_out/gemv_v4dbg_$qt.err; echo "v4 stream-only $qt rc=$?"
done
python - <<'P'
import json
for f in ["gemv_v4_nf4","gemv_old_nf4","gemv_v4dbg_nf4","gemv_v4_int4","gemv_old_int4","gemv_v4dbg_int4"]:
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().split("\n")[-1])
        print(f, d["per_layer_us"], d["step_est_us"], [(r["shape"][:10], r["chain_us"], r["chain_GBs"], r["cold_us"]) for r in d["rows"]])
    except Exception as e: print(f, "ERR", e)
P
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench_default.json
timeout 900 python bench.py --no-cpu-baseline --steps 256 --warmup 5 > gpurun_out/bench_256.json 2> gpurun_out/bench_256.err; echo "bench256 rc=$?"; cut -c1-400 gpurun_out/bench_256.json
./tools/micro/launch_floor > gpurun_out/launch_floor.txt 2>&1; cat gpurun_out/launch_floor.txt
