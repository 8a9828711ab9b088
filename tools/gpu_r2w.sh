#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 300 python -u -m pytest tests/test_gpu_model.py tests/test_gpu_checkpoint.py -x -q -m gpu --timeout 180 -k "single_resident or checkpoint or sampling" > gpurun_out/pytest_sc.log 2>&1; echo "rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_sc.log | grep -E "passed|failed|Error|assert" | tail -6
