#!/bin/bash
# usage: gpu_tpN.sh N workload [env...]
set -u
mkdir -p gpurun_out
N=$1; WL=$2
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --no-cpu-baseline --steps 128 --warmup 8 --workload $WL > gpurun_out/bench_tp${N}_$WL.json 2> gpurun_out/bench_tp${N}_$WL.err; echo "bench tp$N $WL rc=$?"; cut -c1-330 gpurun_out/bench_tp${N}_$WL.json; grep -v "^\*\*\*\|OMP_NUM\|NCCL version" gpurun_out/bench_tp${N}_$WL.err | tail -5
