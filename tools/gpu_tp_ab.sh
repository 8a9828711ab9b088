#!/bin/bash
# usage: gpu_tp_ab.sh N "MIN_A MIN_B" "workloads": tensor-parallel bench with the two-stage exchange threshold at MIN_A / MIN_B
set -u
N=$1
mkdir -p gpurun_out/records
for WL in $3; do
for MIN in $2; do
  B200_TP_TWO_STAGE_MIN=$MIN timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --no-cpu-baseline --steps 128 --warmup 8 --workload $WL > gpurun_out/records/tp${N}_${WL}_min$MIN.json 2> gpurun_out/records/tp${N}_${WL}_min$MIN.err; echo "tp$N $WL two-stage-min $MIN rc=$?"
  python - <<P
import json
try:
    d=json.loads(open("gpurun_out/records/tp${N}_${WL}_min$MIN.json").read().strip().split("\n")[-1]); print("   ", round(d["value"],1), "tok/s", round(d["ms_per_step"],3), "ms")
except Exception as e: print("ERR", e)
P
done
done
