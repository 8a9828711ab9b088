#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 200 python -u -m pytest tests/test_gpu_kernels.py -x -q -m gpu --timeout 40 -k "fused_attention" > gpurun_out/pytest_fa.log 2>&1; rc=$?; echo "fused attn rc=$rc"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_fa.log | grep -E "passed|failed|Error|assert|Timeout" | tail -6
if [ $rc -ne 0 ]; then exit 0; fi
timeout 300 python -u -m pytest tests/test_gpu_model.py -x -q -m gpu --timeout 120 -k "toy or llama7b or chat or sampling" > gpurun_out/pytest_m.log 2>&1; echo "model rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_m.log | grep -E "passed|failed|Error|assert" | tail -4
for mx in 1024 0; do
B200_FUSE_ATTN_MAX_SEQ=$mx timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_fa$mx.json 2> gpurun_out/bench_fa$mx.err; python - <<P
import json
d=json.loads(open("gpurun_out/bench_fa$mx.json").read().strip().split("\n")[-1])
print("fuse max_seq $mx", round(d["value"],1), round(d["ms_per_step"],4), round(d["e2e"]["value"],1), d["launches_per_step"])
P
done
