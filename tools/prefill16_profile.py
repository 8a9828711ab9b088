#!/usr/bin/env python
"""Development tool: kernel-level breakdown of the 16-token prompt prefill that generate() runs before the decode loop."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from lit_gpt_b200 import QuantPrecision
from lit_gpt_b200.fabric import Fabric
fabric = Fabric(devices=1, plugins=QuantPrecision("nf4", torch.bfloat16))
model, cfg = bench.build_random_model("Llama-2-7b-hf", "nf4", 64, 1, fabric)
dev = fabric.device
x = torch.randint(0, 32000, (1, 16), device=dev); pos = torch.arange(16, device=dev)
with torch.inference_mode():
    for _ in range(3): model(x, pos)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); model(x, pos); e1.record(); torch.cuda.synchronize()
    print("prefill(16) ms:", e0.elapsed_time(e1))
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        model(x, pos); torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=14, max_name_column_width=60))
