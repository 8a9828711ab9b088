#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_noalloc_alone.log | tail -4
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_gpu.log | tail -8
timeout 300 python tools/gemv_bench.py --qt nf4 --l2pf 0 --tiny > gpurun_out/gemv_nf4_v4.log 2>&1; grep shape gpurun_out/gemv_nf4_v4.log | cut -c1-200; tail -1 gpurun_out/gemv_nf4_v4.log | cut -c1-160
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/bench_default.json
