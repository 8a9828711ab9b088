#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "tiled" > gpurun_out/pytest_tiled.log 2>&1; echo "tiled rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_tiled.log | tail -8
timeout 900 python -m pytest tests/test_gpu_tp_shared.py -x -q -m gpu -s > gpurun_out/pytest_tp_shared.log 2>&1; echo "tp_shared rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_tp_shared.log | grep -v "^$" | tail -24
timeout 900 python -m pytest tests/test_gpu_model.py tests/test_gpu_checkpoint.py -x -q -m gpu > gpurun_out/pytest_model.log 2>&1; echo "model rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_model.log | tail -5
