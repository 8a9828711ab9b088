#!/usr/bin/env python
"""Per-shape timing of the fused decode attention (RoPE + append + split-K) through the C ABI (development tool).

    python tools/attn_bench.py [--splits 0]

Each shape runs as a CUDA graph over `copies` distinct KV caches (> L2 in total), the way the kernel runs inside a decode
step.  GB/s = filled K + V rows (GQA-native, bf16) / time.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

SHAPES = [  # name, H, G, ctx, B
    ("7b ctx27", 32, 32, 27, 1), ("7b ctx60", 32, 32, 60, 1), ("7b ctx100", 32, 32, 100, 1), ("7b ctx112", 32, 32, 112, 1), ("7b ctx176", 32, 32, 176, 1), ("7b ctx272", 32, 32, 272, 1), ("7b ctx2048", 32, 32, 2048, 1), ("13b ctx2048", 40, 40, 2048, 1), ("13b ctx2048 b4", 40, 40, 2048, 4),
    ("13b ctx2048 b16", 40, 40, 2048, 16), ("mistral ctx4096", 32, 8, 4096, 1), ("70b tp2 ctx2048", 32, 4, 2048, 1), ("70b tp8 ctx2048", 8, 1, 2048, 1),
]


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--splits", type=int, default=0)
    ap.add_argument("--reps", type=int, default=20)
    args = ap.parse_args()
    from lit_gpt_b200 import kernels as K
    from oracle import model_ref as M

    dev = torch.device("cuda")
    hs = 128
    for name, H, G, ctx, B in SHAPES:
        max_seq = ctx + 16
        kv_bytes = 2 * B * G * ctx * hs * 2
        copies = max(2, int(300e6 // kv_bytes) + 1)
        caches = [(torch.randn(B, G, max_seq, hs, device=dev).bfloat16(), torch.randn(B, G, max_seq, hs, device=dev).bfloat16()) for _ in range(copies)]
        qkv = torch.randn(B, 1, (H + 2 * G) * hs, device=dev).bfloat16()
        cos, sin = M.build_rope_cache(max_seq, hs, pos_dtype=torch.float32)
        cos, sin = cos.to(dev), sin.to(dev)
        pos = torch.tensor([ctx], device=dev)

        def run(i):
            return K.attn_decode(qkv, cos, sin, pos, caches[i][0], caches[i][1], H, G, hs, hs, num_splits=args.splits)

        run(0)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for i in range(copies):
                run(i)
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.reps):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / (args.reps * copies)
        print(json.dumps({"shape": name, "H": H, "G": G, "ctx": ctx, "B": B, "kv_MB": round(kv_bytes / 1e6, 2), "us": round(us, 2),
                          "GBs": round(kv_bytes / us / 1e3, 1), "copies": copies}), flush=True)
        del caches


if __name__ == "__main__":
    main()
