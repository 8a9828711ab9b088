#!/bin/bash
set -u
mkdir -p gpurun_out
N=${1:-8}
run() { # name, extra args
  local name=$1; shift
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --no-cpu-baseline "$@" > gpurun_out/bench_tp${N}_$name.json 2> gpurun_out/bench_tp${N}_$name.err; echo "bench tp$N $name rc=$?"; cut -c1-330 gpurun_out/bench_tp${N}_$name.json; grep -v "^\*\*\*\|OMP_NUM\|NCCL version" gpurun_out/bench_tp${N}_$name.err | tail -3
}
run 7b --steps 128 --warmup 8
run 70b --steps 128 --warmup 8 --workload llama2-70b-int4-bs1-ctx2048
B200_TP_NCCL=1 run 70b_nccl --steps 128 --warmup 8 --workload llama2-70b-int4-bs1-ctx2048
