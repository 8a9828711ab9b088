#!/usr/bin/env python
"""Per-shape timing of the bnb.int8 Linear (b200_linear_int8: 3 activation kernels + the int8 tensor-core product)
through the C ABI (development tool).  Decode rows are timed as a CUDA graph over weight copies larger than L2;
GB/s = N * K weight bytes / time.  Prefill rows report int8 TOP/s.

    python tools/int8_bench.py [--model Llama-2-7b-hf] [--batch 1] [--prefill 2048]
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="Llama-2-7b-hf")
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--prefill", type=int, default=2048)
    ap.add_argument("--reps", type=int, default=20)
    args = ap.parse_args()
    from lit_gpt_b200 import Config
    from lit_gpt_b200 import kernels as K

    c = Config.from_name(args.model)
    d, I = c.n_embd, c.intermediate_size
    qkv = (c.n_head + 2 * c.n_query_groups) * c.head_size
    shapes = [("attn.attn", qkv, d), ("attn.proj", d, d), ("mlp.fc_1", I, d), ("mlp.proj", d, I)]
    dev = torch.device("cuda")
    rows = []
    for name, N, Kd in shapes:
        copies = max(2, int(300e6 // (N * Kd)) + 1)
        mats = []
        for i in range(copies):
            w = torch.empty(N, Kd, dtype=torch.float16, device=dev).normal_(0, 0.02)
            mats.append(K.quantize_int8(w))
            del w
        x = torch.randn(args.batch, Kd, device=dev).clamp_(-5, 5).half()
        x[0, 5] = 7.0  # one outlier column, as real activations have a few
        K.linear_int8(x, *mats[0])
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            with torch.cuda.graph(g, stream=s):
                for i in range(copies):
                    K.linear_int8(x, *mats[i])
        for _ in range(3):
            g.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.reps):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / (args.reps * copies)
        row = {"shape": name, "N": N, "K": Kd, "MB": round(N * Kd / 1e6, 2), "decode_us": round(us, 2), "decode_GBs": round(N * Kd / us / 1e3, 1)}
        if args.prefill:
            xp = torch.randn(args.prefill, Kd, device=dev).clamp_(-5, 5).half()
            K.linear_int8(xp, *mats[0])
            torch.cuda.synchronize()
            e0.record()
            for _ in range(5):
                K.linear_int8(xp, *mats[0])
            e1.record()
            torch.cuda.synchronize()
            pus = e0.elapsed_time(e1) * 1e3 / 5
            row.update(prefill_us=round(pus, 1), prefill_TOPs=round(2 * args.prefill * N * Kd / pus / 1e6, 1))
        rows.append(row)
        del mats
    print(json.dumps({"model": args.model, "batch": args.batch, "prefill_rows": args.prefill, "rows": rows}))


if __name__ == "__main__":
    main()
