#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 400 python -u -m pytest tests/test_gpu_tp_shared.py -x -q -m gpu -s --timeout 200 -k "lm_head" > gpurun_out/pytest_lmh.log 2>&1; echo "lm_head sharded rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_lmh.log | grep -E "^\[tp|passed|failed|Error|assert" | tail -12
