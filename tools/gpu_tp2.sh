#!/bin/bash
set -u
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_model.py -m gpu -x -q -k tp2 > gpurun_out/pytest_tp2.log 2>&1; echo "pytest tp2 rc=$?"; tail -15 gpurun_out/pytest_tp2.log
for mode in 0 1; do
  B200_TP_NCCL=$mode timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 128 --warmup 8 --no-cpu-baseline > gpurun_out/bench_tp2_nccl$mode.json 2> gpurun_out/bench_tp2_nccl$mode.err; echo "bench tp2 nccl=$mode rc=$?"; cat gpurun_out/bench_tp2_nccl$mode.json; tail -3 gpurun_out/bench_tp2_nccl$mode.err
done
