#!/usr/bin/env python
"""Timing of the persistent decode-step kernel on synthetic Llama-shaped programs (development tool).

    python tools/step_bench.py [--model Llama-2-7b-hf] [--qt nf4] [--batch 1] [--layers 32] [--ctx 272]
                               [--lookahead 0,256,1024] [--variants full,noattn,tiny]

variants: full = embed + blocks + lm_head; noattn = the same without the ATTN ops (GEMV chain + barriers only);
tiny = the same op count with 64-row matrices (barrier / prologue cost only).
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
    ap.add_argument("--layers", type=int, default=32)
    ap.add_argument("--ctx", type=int, default=272)
    ap.add_argument("--lookahead", default="0")
    ap.add_argument("--variants", default="full,noattn,tiny")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--trace", action="store_true")
    ap.add_argument("--debug", default="0")
    args = ap.parse_args()

    from lit_gpt_b200 import Config
    from lit_gpt_b200 import kernels as K

    c = Config.from_name(args.model)
    d, I, V, H, G, hs = c.n_embd, c.intermediate_size, c.padded_vocab_size, c.n_head, c.n_query_groups, c.head_size
    qkv_n = (H + 2 * G) * hs
    dev = torch.device("cuda")
    B, qt = args.batch, args.qt
    bs = 128 if qt == "int4" else 64
    bpw = 0.5 + (4 / 128 if qt == "int4" else 4 / 64)

    def tiled(N, Kd):
        return torch.randint(0, 255, (K.tiled_bytes(N, Kd, qt),), dtype=torch.uint8, device=dev)

    def new(n):
        return torch.zeros(B, n, dtype=torch.bfloat16, device=dev)

    max_seq = args.ctx + 16
    cos = torch.ones(max_seq, hs, device=dev)
    sin = torch.zeros(max_seq, hs, device=dev)
    pos = torch.tensor([args.ctx], device=dev)
    idx = torch.zeros(B, dtype=torch.int64, device=dev)
    wte = torch.zeros(V, d, dtype=torch.bfloat16, device=dev)
    nw = torch.ones(d, dtype=torch.bfloat16, device=dev)

    def build(variant):
        P = K.DecodeProgram(B, qt, bs, dev)
        tiny = variant == "tiny"
        nbytes = 0
        x = new(d)
        P.embed(idx, wte, x)
        for _ in range(args.layers):
            n1 = 64 if tiny else qkv_n
            qkv = new(qkv_n)
            P.gemv(x, tiled(n1, d), n1, d, qkv[:, :n1].contiguous() if tiny else qkv, norm_weight=nw)
            y = new(H * hs)
            if variant == "full":
                kc = torch.zeros(B, G, max_seq, hs, dtype=torch.bfloat16, device=dev)
                P.attn(qkv, cos, sin, pos, kc, torch.zeros_like(kc), y, H, G, hs, hs)
            if tiny:
                a, b2, c2 = new(64), new(64), new(64)
                P.gemv(y, tiled(64, d), 64, d, a)
                P.gemv(x, tiled(64, d), 64, d, b2, tiled2=tiled(64, d), norm_weight=nw)
                P.gemv(new(I), tiled(64, I), 64, I, c2)
                nbytes += 4 * 64 * d * bpw + 64 * I * bpw
            else:
                x2 = new(d)
                P.gemv(y, tiled(d, d), d, d, x2, residual=x)
                act = new(I)
                P.gemv(x2, tiled(I, d), I, d, act, tiled2=tiled(I, d), norm_weight=nw)
                x3 = new(d)
                P.gemv(act, tiled(d, I), d, I, x3, residual=x2)
                x = x3
                nbytes += (qkv_n * d + d * d + 3 * d * I) * bpw
        if not tiny:
            P.gemv(x, tiled(V, d), V, d, new(V), norm_weight=nw)
            nbytes += V * d * bpw
        return P.finalize(), nbytes

    out = []
    for variant in args.variants.split(","):
        P, nbytes = build(variant)
        for la, dbg in [(int(v), int(d)) for v in args.lookahead.split(",") for d in args.debug.split(",")]:
            K.decode_step_tune(la, False, dbg)
            P.run()
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                P.run()
            g.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(args.reps):
                g.replay()
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / args.reps
            row = {"variant": variant, "lookahead_kb": la, "debug": dbg, "us_per_step": round(us, 1), "GBs": round(nbytes / us / 1e3, 1), "ops": len(P._ops)}
            print(json.dumps(row), flush=True)
            out.append(row)
            if args.trace:
                K.decode_step_tune(la, True, dbg)
                P.run()
                torch.cuda.synchronize()
                K.decode_step_tune(la, False, dbg)
                t = P.trace().double() / 1.965e3  # us at 1965 MHz
                n = t.shape[0]
                kinds = [o.kind for o in P._ops]
                names = {0: "gemv", 1: "attn", 2: "embed"}
                # per op: prologue, main, arrive (bar+fence+atomic), wait (spin), handoff (to next op start)
                rows = []
                for i in range(n - 1):
                    pro = t[i, 1] - t[i, 0] if kinds[i] == 0 else 0.0
                    main = t[i, 2] - (t[i, 1] if kinds[i] == 0 else t[i, 0])
                    rows.append((i, kinds[i], float(pro), float(main), float(t[i, 3] - t[i, 2]), float(t[i, 4] - t[i, 3]), float(t[i + 1, 0] - t[i, 4]), float(t[i, 5]), float(t[i, 6]), float(t[i, 7])))
                import collections
                agg = collections.defaultdict(list)
                for i, k, *v in rows:
                    slot = "embed" if k == 2 else ("attn" if k == 1 else f"gemv{(i - 1) % (5 if variant == 'full' else 4)}")
                    agg[slot].append(v)
                for slot, vs in sorted(agg.items()):
                    m = torch.tensor(vs).mean(0).tolist()
                    print(f"   trace[{variant}] {slot:7s} n={len(vs):3d} prologue={m[0]:6.2f} main={m[1]:6.2f} arrive={m[2]:5.2f} wait={m[3]:6.2f} handoff={m[4]:5.2f} | main split: data-wait={m[5]:5.2f} compute={m[6]:5.2f} reduce={m[7]:5.2f} us (CTA 0 warp 0)", flush=True)
                print(f"   trace total {float(t[n - 1, 2] - t[0, 0]):.1f} us", flush=True)
        del P
        torch.cuda.empty_cache()
    K.decode_step_tune(0)


if __name__ == "__main__":
    main()
