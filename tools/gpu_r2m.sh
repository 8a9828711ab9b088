#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -u -m pytest tests/test_gpu_model.py -x -q -m gpu --timeout 180 -s -k "long_context or chat_stream" > gpurun_out/pytest_new.log 2>&1; echo "pytest rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_new.log | grep -E "^\[|passed|failed|Error|assert|Timeout" | tail -20
