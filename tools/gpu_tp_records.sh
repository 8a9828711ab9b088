#!/bin/bash
# usage: gpu_tp_records.sh N : tensor-parallel records at world N (7B nf4 headline workload + 70B int4), plus the multi-GPU pytest at N = 2
set -u
N=$1
mkdir -p gpurun_out/records
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
if [ "$N" = "2" ]; then
  timeout 900 python -m pytest tests/test_gpu_model.py -x -q -m gpu -k "tp2" -s > gpurun_out/records/pytest_tp2_multi_gpu.log 2>&1; echo "pytest tp2 rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/records/pytest_tp2_multi_gpu.log | grep -v "^$" | tail -8
fi
for WL in llama2-7b-nf4-bs1 llama2-70b-int4-bs1-ctx2048; do
  timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --no-cpu-baseline --steps 128 --warmup 8 --workload $WL > gpurun_out/records/tp${N}_$WL.json 2> gpurun_out/records/tp${N}_$WL.err; echo "bench tp$N $WL rc=$?"; cut -c1-200 gpurun_out/records/tp${N}_$WL.json; grep -v "^\*\*\*\|OMP_NUM\|NCCL version\|^DEBUG" gpurun_out/records/tp${N}_$WL.err | tail -4
done
python - <<'P'
import json,glob
for f in sorted(glob.glob("gpurun_out/records/tp*.json")):
    try:
        d=json.loads(open(f).read().strip().split("\n")[-1])
        print(f.split("/")[-1], round(d["value"],1), "tok/s", round(d["ms_per_step"],3), "ms", "gemv frac", round(d["roofline"]["frac"],3), "step frac", round(d["roofline"]["whole_step_frac"],3), d["clocks"].get("sm_mhz"), d["clocks"].get("samples"))
    except Exception as e: print(f, "ERR", e)
P
