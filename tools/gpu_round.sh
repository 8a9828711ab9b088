#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_model.py -m gpu -x -q -k "generic or qgemv or tp2" > gpurun_out/pytest_generic.log 2>&1; echo "pytest rc=$?"; grep -v "^DEBUG\|Deprec\|warnings.warn" gpurun_out/pytest_generic.log | tail -5
python - <<'PY'
import torch, sys
sys.path.insert(0,'.')
from lit_gpt_b200 import kernels as K
w=torch.randn(4096,1376,device='cuda').bfloat16()*0.02
p,a=K.quantize_4bit(w,'nf4')
x=torch.randn(1,1376,device='cuda').bfloat16()
y=K.qgemv(x,p,a,4096,1376,'nf4',64)
torch.cuda.synchronize()
e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(50): K.qgemv(x,p,a,4096,1376,'nf4',64)
e1.record(); torch.cuda.synchronize()
print("generic 4096x1376 us:", e0.elapsed_time(e1)*1e3/50)
PY
