#!/bin/bash
set -u
mkdir -p gpurun_out
B200_TP_TWO_STAGE_MIN=2 timeout 400 python -u -m pytest tests/test_gpu_tp_shared.py -x -q -m gpu -s --timeout 300 > gpurun_out/pytest_tp_shared.log 2>&1; echo "tp_shared (two-stage from world 2) rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_tp_shared.log | grep -E "^\[tp|passed|failed|Error" | tail -8
