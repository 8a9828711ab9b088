#!/usr/bin/env python
"""Per-shape timing of the fused dequant-GEMV through the C ABI (development tool, not the headline bench).

    python tools/gemv_bench.py [--model Llama-2-7b-hf] [--qt nf4] [--batch 1] [--ks 0] [--occ 0] [--reps 50]

Each shape is timed two ways with CUDA events on the launching stream:
  cold  : a > L2-sized buffer is rewritten between launches (every weight byte comes from HBM),
  chain : `copies` distinct weight copies (> L2 in total) launched back to back in one CUDA graph, which is how
          the kernel runs inside a decode step (PDL overlap between consecutive launches included).
GB/s = algorithmic bytes (packed weights + scales) / time.
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
    ap.add_argument("--qt", default="nf4")
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--ks", type=int, default=0)
    ap.add_argument("--occ", type=int, default=0)
    ap.add_argument("--reps", type=int, default=30)
    ap.add_argument("--pdl", type=int, default=1)
    ap.add_argument("--json", action="store_true")
    ap.add_argument("--kernel", choices=["tiled", "ref"], default="tiled")
    ap.add_argument("--depth", type=int, default=0)
    ap.add_argument("--l2pf", type=int, default=1)
    ap.add_argument("--tiny", action="store_true")
    ap.add_argument("--only", default="", help="substring of the shape names to run (profiling)")
    args = ap.parse_args()

    from lit_gpt_b200 import Config
    from lit_gpt_b200 import kernels as K

    c = Config.from_name(args.model)
    d, I, V = c.n_embd, c.intermediate_size, c.padded_vocab_size
    qkv = (c.n_head + 2 * c.n_query_groups) * c.head_size
    shapes = [("attn.attn+norm", qkv, d, "norm"), ("attn.proj+res", d, d, "res"), ("fc_1|fc_2 swiglu+norm", I, d, "swiglu"),
              ("mlp.proj+res", d, I, "res"), ("lm_head+norm", V, d, "norm")]
    if args.tiny:  # one tile per SM: the time of such a launch is the fixed cost of a kernel boundary
        shapes += [("tiny+norm", 148 * 16, d, "norm"), ("tiny+res", 148 * 16, d, "res")]
    if args.only:
        shapes = [s for s in shapes if args.only in s[0]]
    dev = torch.device("cuda")
    K.set_pdl(bool(args.pdl))
    K.qgemv_tune(args.ks, args.occ)
    K.qgemv_tiled_tune(args.depth, args.occ, bool(args.l2pf))
    tiled = args.kernel == "tiled"
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
    rows = []
    for name, N, Kd, kind in shapes:
        nmat = 2 if kind == "swiglu" else 1
        bytes_alg = nmat * N * Kd * (0.5 + (4 / 64 if args.qt != "int4" else 4 / 128))
        copies = min(64, max(2, int(300e6 // bytes_alg) + 1))
        mats = []
        for i in range(copies * nmat):
            w = torch.empty(N, Kd, dtype=torch.bfloat16, device=dev).normal_(0, 0.02)
            if args.qt == "int4":
                qw, _, _, sz = K.quantize_int4(w, 128)
                mats.append((K.tile_4bit(qw, sz, N, Kd, "int4", 128), None, 128) if tiled else (qw, sz, 128))
            else:
                p, a = K.quantize_4bit(w, args.qt)
                mats.append((K.tile_4bit(p, a, N, Kd, args.qt, 64), None, 64) if tiled else (p, a, 64))
            del w
        x = torch.randn(args.batch, Kd, device=dev).to(torch.bfloat16)
        nw = torch.ones(Kd, dtype=torch.bfloat16, device=dev)
        res = torch.randn(args.batch, N, device=dev).to(torch.bfloat16)
        y = torch.empty(args.batch, N, dtype=torch.bfloat16, device=dev)

        def run(i: int) -> None:
            p, s, bs = mats[i * nmat]
            kw = {}
            if kind in ("norm", "swiglu") and args.batch <= 8:   # above 8 rows the model normalises in its own launch (GPT._fused_decode)
                kw.update(norm_weight=nw, norm_eps=1e-5)
            if kind == "res":
                kw.update(residual=res)
            if tiled:
                if kind == "swiglu":
                    kw.update(tiled2=mats[i * nmat + 1][0])
                nxt = mats[((i + 1) % copies) * nmat][0] if args.l2pf else None  # the chain's next launch reads these
                K.qgemv_tiled(x, p, N, Kd, args.qt, bs, out=y, prefetch=nxt, **kw)
                return
            if kind == "swiglu":
                kw.update(packed2=mats[i * nmat + 1][0], scale2=mats[i * nmat + 1][1])
            K.qgemv(x, p, s, N, Kd, args.qt, bs, out=y, **kw)

        run(0)
        torch.cuda.synchronize()
        # cold: flush L2 between launches
        ts = []
        for r in range(min(args.reps, 12)):
            flush.fill_(r & 0xFF)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            run(r % copies)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        cold_us = sorted(ts)[len(ts) // 2]
        # chain: graph of `copies` launches over distinct weights
        g = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            run(0)
        torch.cuda.current_stream().wait_stream(side)
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
        chain_us = e0.elapsed_time(e1) * 1e3 / (args.reps * copies)
        row = {"shape": name, "N": N, "K": Kd, "MB": round(bytes_alg / 1e6, 2), "cold_us": round(cold_us, 2),
               "cold_GBs": round(bytes_alg / cold_us / 1e3, 1), "chain_us": round(chain_us, 2),
               "chain_GBs": round(bytes_alg / chain_us / 1e3, 1), "copies": copies}
        rows.append(row)
        if not args.json:
            print(row, flush=True)
        del mats
        torch.cuda.empty_cache()
    if args.only:
        print(json.dumps(rows), flush=True)
        return
    per_layer = sum(r["chain_us"] for r in rows[:4])
    summary = {"model": args.model, "qt": args.qt, "batch": args.batch, "kernel": args.kernel, "depth": args.depth, "l2pf": args.l2pf, "ks": args.ks, "occ": args.occ, "pdl": args.pdl,
               "per_layer_us": round(per_layer, 2), "step_est_us": round(per_layer * c.n_layer + rows[4]["chain_us"], 1), "rows": rows}
    print(json.dumps(summary), flush=True)


if __name__ == "__main__":
    main()
