#!/usr/bin/env python
"""Development tool: where does a dequant-GEMV launch of the decode chain spend its time?

    python tools/trace_step.py [--layers 4] [--qt nf4] [--model Llama-2-7b-hf]

Captures one fused decode step into a CUDA graph with b200_qgemv_tiled_trace armed, replays it, and prints per launch
(median over CTAs, ns relative to the launch's first CTA entry): entry skew, dependency release, prologue end, first
stage arrival, last stage consumed, last output stored; plus the gap between consecutive launches."""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="Llama-2-7b-hf")
    ap.add_argument("--qt", default="nf4")
    ap.add_argument("--layers", type=int, default=4)
    ap.add_argument("--ctx", type=int, default=256)
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--dump", type=int, default=-1, help="launch slot whose slowest and fastest CTAs are listed stamp by stamp")
    ap.add_argument("--prologue", action="store_true", help="cycle stamps inside the prologue instead of the stream phases")
    args = ap.parse_args()
    import bench
    from lit_gpt_b200 import QuantPrecision, _cabi as C
    from lit_gpt_b200.fabric import Fabric

    fabric = Fabric(devices=1, plugins=QuantPrecision(args.qt, torch.bfloat16))
    model, _ = bench.build_random_model(args.model, args.qt, args.ctx + 32, args.batch, fabric, n_layer=args.layers)
    dev = fabric.device
    tok = torch.full((args.batch, 1), 7, device=dev)
    pos = torch.tensor([args.ctx], device=dev)
    with torch.inference_mode():
        model(torch.randint(0, 1000, (args.batch, 16), device=dev), torch.arange(16, device=dev))
        for _ in range(3):
            model(tok, pos)
        torch.cuda.synchronize()
        slots = 4 * args.layers + 1
        buf = torch.zeros(slots * 160 * 8, dtype=torch.int64, device=dev)
        lib = C.load()
        if args.prologue:
            lib.b200_qgemv_tiled_tune(97, 0, 0)
        lib.b200_qgemv_tiled_trace(buf.data_ptr(), slots)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            model(tok, pos)
        lib.b200_qgemv_tiled_trace(None, 0)
        for _ in range(5):
            g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        print(f"step {e0.elapsed_time(e1) / 20 * 1e3:.1f} us for {args.layers} layers + lm_head")
    t = buf.cpu().numpy().reshape(slots, 160, 8).astype(np.float64)
    names = ["qkv", "proj", "fc", "mproj"]
    if args.dump >= 0:
        a = t[args.dump]
        used = a[:, 0] > 0
        t0 = a[used][:, 0].min()
        order = np.argsort(np.where(used, a[:, 5], 0))
        order = [i for i in order if used[i]]
        print("cta | entry dep-release prologue-end 1st-stage last-consumed done producer-done gather-start")
        for i in sorted(order):
            r = a[i] - t0
            print(f"{i:4d} | {r[0]:6.0f} {r[1]:6.0f} {r[2]:6.0f} {r[3]:6.0f} {r[4]:6.0f} {r[5]:6.0f} {r[6]:6.0f} {r[7]:6.0f}")
    if args.prologue:
        print("cycles after the dependency release (median over CTAs): x loaded+ssq | 1st barrier | scaled | prologue end (ns) | first stage passed")
        for s in range(slots):
            a = t[s][t[s][:, 0] > 0]
            if a.size == 0:
                continue
            name = names[s % 4] if s < slots - 1 else "lm_head"
            late = a[:, 0] > a[:, 1].min()  # CTAs that entered after the first release
            print(f"{s:3d} {name:7s} | {np.median(a[:, 3]):6.0f} | {np.median(a[:, 6]):6.0f} | {np.median(a[:, 7]):6.0f} | {np.median(a[:, 2] - a[:, 1]):6.0f} ns | {np.median(a[:, 4]):6.0f} | entered late: {int(late.sum())}")
        return
    print("launch      ctas | entry: med max | dep-release | prologue-end | 1st-stage | producer-done | last-consumed | reducer-done: med max | total | gap-to-next-entry")
    prev_end = None
    for s in range(slots):
        used = t[s][:, 0] > 0
        a = t[s][used]
        if a.size == 0:
            continue
        t0 = a[:, 0].min()
        r = a - t0
        med = lambda k: np.median(r[:, k][a[:, k] > 0]) if (a[:, k] > 0).any() else float("nan")
        mx = lambda k: r[:, k][a[:, k] > 0].max() if (a[:, k] > 0).any() else float("nan")
        name = names[s % 4] if s < slots - 1 else "lm_head"
        gap = "" if prev_end is None else f"{t0 - prev_end:7.0f}"
        end = max(mx(5), mx(7)) if (a[:, 7] > 0).any() and name == "qkv" else mx(5)
        extra = f" | fused attention: counters ready med {med(6):6.0f}, done med {med(7):6.0f} max {mx(7):6.0f}" if name == "qkv" and (a[:, 7] > 0).any() and med(7) > med(5) else ""
        print(f"{s:3d} {name:7s} {int(used.sum()):4d} | {med(0):6.0f} {mx(0):6.0f} | {med(1):6.0f} | {med(2):6.0f} | {med(3):6.0f} | {med(6):6.0f} | {med(4):6.0f} | {med(5):6.0f} {mx(5):6.0f} | {end:6.0f} | prev-end->entry {gap}{extra}")
        prev_end = max(a[:, 5][a[:, 5] > 0].max(), a[:, 7].max() if name == "qkv" and (a[:, 7] > a[:, 5]).any() else 0)


if __name__ == "__main__":
    main()
