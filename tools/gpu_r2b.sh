#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tp_shared.py -x -q -m gpu -s > gpurun_out/pytest_tp_shared.log 2>&1; echo "tp_shared rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_tp_shared.log | tail -30
timeout 600 python tools/trace_step.py --layers 4 > gpurun_out/trace_nf4.txt 2>&1; echo "trace rc=$?"; cat gpurun_out/trace_nf4.txt | tail -25
