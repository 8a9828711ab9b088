#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_model.py -m gpu -x -q -k "traced" > gpurun_out/pytest_trace.log 2>&1; echo "pytest rc=$?"; tail -40 gpurun_out/pytest_trace.log
