#!/usr/bin/env python
"""Prefill GEMM timing through the C ABI: tcgen05 dequant-GEMM vs (dequantise kernel + cuBLAS) vs cuBLAS alone.

    python tools/gemm_bench.py [--model Mistral-7B-v0.1] [--qt int4] [--tokens 4096]
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def timed(fn, reps=10):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / reps


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="Llama-2-7b-hf")
    ap.add_argument("--qt", default="nf4")
    ap.add_argument("--tokens", type=int, default=4096)
    args = ap.parse_args()
    from lit_gpt_b200 import Config
    from lit_gpt_b200 import kernels as K

    c = Config.from_name(args.model)
    d, I, V = c.n_embd, c.intermediate_size, c.padded_vocab_size
    qkv = (c.n_head + 2 * c.n_query_groups) * c.head_size
    M = args.tokens
    rows = []
    for name, N, Kd in [("attn.attn", qkv, d), ("attn.proj", d, d), ("mlp.fc", I, d), ("mlp.proj", d, I)]:
        w = torch.empty(N, Kd, dtype=torch.bfloat16, device="cuda").normal_(0, 0.02)
        x = torch.randn(M, Kd, device="cuda").to(torch.bfloat16)
        if args.qt == "int4":
            qw, scales, qz, sz = K.quantize_int4(w, 128)
            packed, scale, bs = qw, sz, 128
            deq = lambda: K.dequantize_int4(qw, scales, qz, Kd, 128)
        else:
            packed, scale = K.quantize_4bit(w, args.qt)
            bs = 64
            deq = lambda: K.dequantize_4bit(packed, scale, (N, Kd), torch.bfloat16, args.qt)
        wb = deq()
        y = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
        t_k = timed(lambda: K.qgemm(x, packed, scale, N, Kd, args.qt, bs, out=y))
        t_c = timed(lambda: torch.matmul(x, wb.T, out=y))
        t_dc = timed(lambda: torch.matmul(x, deq().T, out=y))
        fl = 2.0 * M * N * Kd
        row = {"layer": name, "M": M, "N": N, "K": Kd, "tcgen05_us": round(t_k, 1), "tcgen05_TFs": round(fl / t_k / 1e6, 1),
               "cublas_us": round(t_c, 1), "cublas_TFs": round(fl / t_c / 1e6, 1), "dequant+cublas_us": round(t_dc, 1),
               "dequant+cublas_TFs": round(fl / t_dc / 1e6, 1)}
        print(json.dumps(row), flush=True)
        rows.append(row)


if __name__ == "__main__":
    main()
