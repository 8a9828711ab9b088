#!/usr/bin/env python
"""Prefill throughput of the quantized model: one `model(x, arange(T))` call through the public
GPT API — tcgen05 dequant-GEMM for every Linear, b200_rope_kv_append + b200_attn_prefill for the attention.

    python tools/prefill_bench.py [--model Mistral-7B-v0.1] [--qt int4] [--tokens 4096] [--reps 5]

Prints one JSON line: prefill tokens/s, algorithmic TFLOP/s (2 * P_lin * T + causal attention, lm_head on all T
positions as the reference computes it, model.py:519) and the attention kernel's own time.
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
    ap.add_argument("--model", default="Mistral-7B-v0.1")
    ap.add_argument("--qt", default="int4")
    ap.add_argument("--tokens", type=int, default=4096)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--layers", type=int, default=None)
    ap.add_argument("--mt3", type=int, default=1, help="0: 256-token GEMM tiles (round 1) instead of 384")
    args = ap.parse_args()

    import bench
    from lit_gpt_b200 import QuantPrecision
    from lit_gpt_b200 import kernels as K
    from lit_gpt_b200.fabric import Fabric

    K.qgemm_tune(bool(args.mt3))
    fabric = Fabric(devices=1, plugins=QuantPrecision(args.qt, torch.bfloat16))
    T = args.tokens
    model, cfg = bench.build_random_model(args.model, args.qt, T, 1, fabric, n_layer=args.layers)
    dev = fabric.device
    g = torch.Generator().manual_seed(1234)
    idx = torch.randint(0, cfg.vocab_size, (1, T), generator=g).to(dev)
    pos = torch.arange(T, device=dev)

    def ev():
        return torch.cuda.Event(enable_timing=True)

    with torch.inference_mode():
        for _ in range(2):
            model(idx, pos)
        torch.cuda.synchronize()
        ts = []
        for _ in range(args.reps):
            e0, e1 = ev(), ev()
            e0.record()
            model(idx, pos)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ms = sorted(ts)[len(ts) // 2]
        # the attention kernel alone, on this model's shapes
        c = model.config
        q = torch.randn(1, c.n_head, T, c.head_size, device=dev).bfloat16()
        cache = model.transformer.h[0].attn.kv_cache
        K.attn_prefill(q, cache.k, cache.v, pos)
        torch.cuda.synchronize()
        e0, e1 = ev(), ev()
        e0.record()
        for _ in range(10):
            K.attn_prefill(q, cache.k, cache.v, pos)
        e1.record()
        torch.cuda.synchronize()
        attn_ms = e0.elapsed_time(e1) / 10
    L = len(model.transformer.h)
    per_layer = cfg.n_embd * (cfg.n_head + 2 * cfg.n_query_groups) * cfg.head_size + cfg.n_embd * cfg.n_embd + 3 * cfg.n_embd * cfg.intermediate_size
    lin_flop = 2.0 * (L * per_layer + cfg.padded_vocab_size * cfg.n_embd) * T
    attn_flop_layer = 4.0 * cfg.n_head * cfg.head_size * T * T / 2
    flop = lin_flop + L * attn_flop_layer
    print(json.dumps({"model": args.model, "qt": args.qt, "tokens": T, "layers": L, "prefill_ms": round(ms, 2),
                      "prefill_tok_s": round(T / ms * 1e3, 1), "algorithmic_tflops": round(flop / ms / 1e9, 1),
                      "attn_kernel_ms_per_layer": round(attn_ms, 3),
                      "attn_kernel_tflops": round(attn_flop_layer / attn_ms / 1e9, 1),
                      "attn_share": round(attn_ms * L / ms, 3)}), flush=True)


if __name__ == "__main__":
    main()
