#!/usr/bin/env python
"""Development tool: QLinear.forward time per call over the number of activation rows (streaming GEMV for <= 16 rows, tcgen05 dequant-GEMM above),
cycling over `--layers` blocks so that the weights come from HBM.   python tools/rows_sweep.py --model Llama-2-7b-hf --qt nf4"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="Llama-2-7b-hf")
    ap.add_argument("--qt", default="nf4")
    ap.add_argument("--layers", type=int, default=8)
    ap.add_argument("--rows", default="1,4,8,16,17,32,64,128")
    args = ap.parse_args()
    import bench
    from lit_gpt_b200 import QuantPrecision
    from lit_gpt_b200.fabric import Fabric

    fabric = Fabric(devices=1, plugins=QuantPrecision(args.qt, torch.bfloat16))
    model, _ = bench.build_random_model(args.model, args.qt, 64, 1, fabric, n_layer=args.layers)
    dev = fabric.device
    names = ["attn.attn", "attn.proj", "mlp.fc_1", "mlp.proj"]
    with torch.inference_mode():
        for name in names:
            mods = [blk.get_submodule(name) for blk in model.transformer.h]
            line = f"{name:10s} ({mods[0].out_features}x{mods[0].in_features})"
            for rows in [int(r) for r in args.rows.split(",")]:
                x = torch.randn(rows, mods[0].in_features, device=dev, dtype=torch.bfloat16)
                for m in mods:
                    m(x)
                torch.cuda.synchronize()
                g = torch.cuda.CUDAGraph()   # device time: the eager loop is bound by ~20 us of Python per call
                with torch.cuda.graph(g):
                    for m in mods:
                        m(x)
                g.replay()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(5):
                    g.replay()
                e1.record()
                torch.cuda.synchronize()
                line += f" | {rows}: {e0.elapsed_time(e1) / (5 * len(mods)) * 1e3:6.1f}"
            print(line + "  us/call", flush=True)


if __name__ == "__main__":
    main()
