#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_checkpoint.py -m gpu -x -q > gpurun_out/pytest_ckpt.log 2>&1; echo "pytest rc=$?"; grep -v "^DEBUG\|Deprecat\|warnings.warn" gpurun_out/pytest_ckpt.log | tail -30
