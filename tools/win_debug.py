import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lit_gpt_b200 import kernels as K
qt, N, Kd, B, depth = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
w = torch.randn(N, Kd, device="cuda").mul_(0.02).to(torch.bfloat16)
if qt == "int4":
    qw, _, _, sz = K.quantize_int4(w, 128); tiled = K.tile_4bit(qw, sz, N, Kd, "int4", 128); bs = 128
else:
    p, a = K.quantize_4bit(w, qt); tiled = K.tile_4bit(p, a, N, Kd, qt, 64); bs = 64
x = torch.randn(B, Kd, device="cuda").to(torch.bfloat16)
K.qgemv_tiled_tune(depth, 0, False)
print("launch", qt, N, Kd, B, depth, flush=True)
y = K.qgemv_tiled(x, tiled, N, Kd, qt, bs)
torch.cuda.synchronize()
print("done", float(y.float().abs().max()), flush=True)
