#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "int4" > gpurun_out/pytest_int4.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_int4.log
timeout 300 python tools/gemv_bench.py --qt int4 --l2pf 0 > gpurun_out/gemv_int4.log 2>&1; tail -1 gpurun_out/gemv_int4.log | cut -c1-200
timeout 300 python tools/gemv_bench.py --qt int4 --l2pf 0 --model Llama-2-13b-hf --batch 4 > gpurun_out/gemv_int4_13b_b4.log 2>&1; tail -1 gpurun_out/gemv_int4_13b_b4.log | cut -c1-1500
