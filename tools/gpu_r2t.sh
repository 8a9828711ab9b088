#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 400 python -u -m pytest tests/test_gpu_model.py -x -q -m gpu --timeout 180 -k "sequential" > gpurun_out/pytest_seq.log 2>&1; echo "seq rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_seq.log | grep -E "passed|failed|Error|assert|skipped" | tail -8
