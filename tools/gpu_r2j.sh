#!/bin/bash
set -u
mkdir -p gpurun_out
for k in "k_windows and nf4 and 1000" "k_windows and int4 and 1000" "k_windows and int4 and 640" "k_windows and int4 and 2048" "k_windows and int4 and 5120-13824-16" "k_windows and nf4 and 5120-13824-16" "k_windows and 5120-13824-9" "k_windows and 96-28672" "k_windows_swiglu" "folded"; do
  timeout 120 python -u -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "$k" > gpurun_out/pytest_k.log 2>&1; rc=$?; echo "[$k] rc=$rc"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_k.log | grep -E "passed|failed|Error|assert" | tail -3
  if [ $rc -eq 124 ]; then nvidia-smi --query-gpu=utilization.gpu --format=csv,noheader; fi
done
