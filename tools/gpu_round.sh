#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "prefill or rope_kv" > gpurun_out/pytest_prefill.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_prefill.log
timeout 900 python -m pytest tests/test_gpu_model.py -m gpu -x -q > gpurun_out/pytest_model.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_model.log
timeout 600 python tools/prefill_bench.py > gpurun_out/prefill_mistral.log 2>&1; tail -2 gpurun_out/prefill_mistral.log
timeout 600 python tools/prefill_bench.py --model Llama-2-7b-hf --qt nf4 > gpurun_out/prefill_llama.log 2>&1; tail -2 gpurun_out/prefill_llama.log
