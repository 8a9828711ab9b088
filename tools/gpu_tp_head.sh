#!/bin/bash
set -u
N=$1
mkdir -p gpurun_out/records
run() { # workload shard
  B200_TP_SHARD_LM_HEAD=$2 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --no-cpu-baseline --steps 128 --warmup 8 --workload $1 > gpurun_out/records/tp${N}_$1_head$2.json 2> gpurun_out/records/tp${N}_$1_head$2.err; echo "tp$N $1 lm_head sharded=$2 rc=$?"
  python - <<P
import json
try:
    d=json.loads(open("gpurun_out/records/tp${N}_$1_head$2.json").read().strip().split("\n")[-1]); print("   ", round(d["value"],1), "tok/s", round(d["ms_per_step"],3), "ms", d["config"].get("lm_head_sharded"))
except Exception as e: print("ERR", e)
P
}
run llama2-7b-nf4-bs1 1
run llama2-7b-nf4-bs1 0
run llama2-70b-int4-bs1-ctx2048 1
