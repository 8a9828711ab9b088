#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_noalloc_alone.log | tail -4
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_gpu.log | tail -8
./tools/micro/launch_floor > gpurun_out/launch_floor.txt 2>&1; cat gpurun_out/launch_floor.txt
