#!/bin/bash
set -u
mkdir -p gpurun_out/records
timeout 900 python -u -m pytest tests/test_gpu_model.py tests/test_gpu_tp_shared.py -q -m gpu -k "tp" -s --timeout 400 > gpurun_out/records/pytest_tp_multi_gpu.log 2>&1; echo "pytest tp rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/records/pytest_tp_multi_gpu.log | grep -E "^\[tp|passed|failed|Error" | tail -14
