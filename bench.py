#!/usr/bin/env python
"""Headline benchmark: Llama-2-7B nf4 weight-only, batch-1 greedy decode tok/s (BASELINE.json configs[1]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

A "step" is one decode token.  `value` = tokens/s over K replays of the captured decode graph with everything
resident in HBM (CUDA events, max over ranks); `e2e` = the same metric through the public `generate()` call with the
prompt in pinned HOST memory and the tokens read back to the host inside the timed region.  N > 1 = the same model
tensor-parallel over N GPUs (strong scaling), launched by torchrun.  Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (config name, quantize mode, batch, prompt tokens, context pre-fill)
    "llama2-7b-nf4-bs1": ("Llama-2-7b-hf", "nf4", 1, 16, 0),
    "llama2-7b-nf4-bs1-ctx2048": ("Llama-2-7b-hf", "nf4", 1, 16, 2032),
    "llama2-13b-int4-bs1-ctx2048": ("Llama-2-13b-hf", "int4", 1, 16, 2032),
    "llama2-13b-int4-bs4-ctx2048": ("Llama-2-13b-hf", "int4", 4, 16, 2032),
    "llama2-13b-int4-bs16-ctx2048": ("Llama-2-13b-hf", "int4", 16, 16, 2032),
    "mistral-7b-int4-bs1-ctx3936": ("Mistral-7B-v0.1", "int4", 1, 16, 3920),  # block_size 4096 bounds prompt + steps
    "mistral-7b-int4-bs1": ("Mistral-7B-v0.1", "int4", 1, 16, 0),
    "llama2-70b-int4-bs1-ctx2048": ("Llama-2-70b-hf", "int4", 1, 16, 2032),
}
# algorithmic bytes per parameter: nf4 / fp4 = 4 bits + one fp32 absmax per 64; int4 g128 = 4 bits + one fp16 scale and
# one 4-bit zero-point per 128 (0.51953).  The kernels stream the fused half2 {scale, zero} form of int4 (0.53125): reported as
# `streamed_bytes_per_param`, not used for the roofline fraction.
BPP = {"nf4": 0.5 + 4 / 64, "fp4": 0.5 + 4 / 64, "int4": 0.5 + 2.5 / 128}
BPP_STREAMED = {"nf4": 0.5 + 4 / 64, "fp4": 0.5 + 4 / 64, "int4": 0.5 + 4 / 128}


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons sampled in-process through NVML every ~2 ms (B200_PROFILING.md clocks line); only the
    samples taken between mark_start() and mark_end() -- the timed region -- are reported."""

    def __init__(self, device: "torch.device"):
        super().__init__(daemon=True)
        self.rows, self._stop_evt, self.t0, self.t1, self.handle, self.nv = [], threading.Event(), None, None, None, None
        try:
            import pynvml as nv

            nv.nvmlInit()
            uuid = str(torch.cuda.get_device_properties(device).uuid)
            try:
                self.handle = nv.nvmlDeviceGetHandleByUUID(("GPU-" + uuid).encode())
            except Exception:
                self.handle = nv.nvmlDeviceGetHandleByIndex(device.index or 0)
            self.nv = nv
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(self.handle, nv.NVML_CLOCK_SM))
        except Exception as e:  # noqa: BLE001
            self.err = str(e)

    def run(self):
        nv = self.nv
        if nv is None:
            return
        while not self._stop_evt.is_set():
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM)
                try:
                    reasons = nv.nvmlDeviceGetCurrentClocksEventReasons(self.handle)
                except Exception:
                    reasons = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
                self.rows.append((time.perf_counter(), float(mhz), int(reasons)))
            except Exception:
                pass
            time.sleep(0.002)

    def mark_start(self):
        self.t0 = time.perf_counter()

    def mark_end(self):
        self.t1 = time.perf_counter()

    def stop(self) -> dict:
        self._stop_evt.set()
        self.join(timeout=1.0)
        nv = self.nv
        if nv is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "source": "nvml unavailable: " + getattr(self, "err", "?")}
        rows = [r for r in self.rows if self.t0 is not None and self.t0 <= r[0] <= (self.t1 or 1e99)]
        bits = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4,
                "hw_power_brake_slowdown": 0x80}
        mask = 0
        for r in rows:
            mask |= r[2]
        return {"sm_mhz": statistics.median([r[1] for r in rows]) if rows else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(n for n, b in bits.items() if mask & b), "samples": len(rows),
                "source": "NVML in-process, 2 ms period, samples inside the timed region only"}


def build_random_model(cfg_name: str, mode: str, max_seq: int, batch: int, fabric, n_layer=None):
    """Random-init weights of the named architecture (N(0, 0.02), seed 1234, bf16), materialised one matrix at a
    time on the GPU, sharded in full precision (generate/tp.py order) and then quantized by the library."""
    from lit_gpt_b200 import GPT, Config, QLinear
    from lit_gpt_b200.generate.tp import shard_lm_head, tensor_parallel

    kw = {} if n_layer is None else {"n_layer": n_layer}
    cfg = Config.from_name(cfg_name, **kw)
    full = Config.from_name(cfg_name, **kw)
    with fabric._precision.module_init_context(), torch.device("meta"):
        model = GPT(cfg)
    full_shapes = {n: (m.out_features, m.in_features) for n, m in model.named_modules() if isinstance(m, QLinear)}
    styles = {}
    if fabric.world_size > 1:
        model = tensor_parallel(fabric, model)
        if os.environ.get("B200_TP_SHARD_LM_HEAD", "1") == "1":   # §8f-3: vocabulary slices per rank, all-gathered over peer memory
            model = shard_lm_head(fabric, model)
        for n in full_shapes:
            styles[n] = 0 if n.endswith(("attn.attn", "fc_1", "fc_2", "fc")) else (1 if n.endswith("proj") else None)
        if getattr(model, "lm_head_sharded", False):
            styles["lm_head"] = 0
    dev = fabric.device
    gen = torch.Generator(device=dev)
    for i, (name, mod) in enumerate(model.named_modules()):
        if isinstance(mod, QLinear):
            gen.manual_seed(1234 + i)
            w = torch.empty(full_shapes[name], dtype=torch.bfloat16, device=dev).normal_(0.0, 0.02, generator=gen)
            if styles.get(name) is not None:
                w = torch.tensor_split(w, fabric.world_size, dim=styles[name])[fabric.global_rank].contiguous()
            mod.weight = torch.nn.Parameter(w, requires_grad=False)
            mod.quantize_()
            del w
        elif isinstance(mod, torch.nn.Embedding):
            gen.manual_seed(1234 + i)
            mod.weight = torch.nn.Parameter(torch.empty(mod.weight.shape, dtype=torch.bfloat16, device=dev).normal_(0.0, 0.02, generator=gen), requires_grad=False)
        elif type(mod).__name__ == "RMSNorm":
            mod.weight = torch.nn.Parameter(torch.ones(mod.weight.shape, dtype=torch.bfloat16, device=dev), requires_grad=False)
    with fabric.init_tensor():
        model.max_seq_length = max_seq
        model.cos, model.sin = model.rope_cache()
        model.set_kv_cache(batch_size=batch)
    model.eval()
    return model, full


def step_bytes(full_cfg, mode: str, ctx: int, batch: int, tp: int, head_sharded: bool = False) -> dict:
    """Algorithmic HBM bytes of one decode step per GPU: packed weights + scales of every Linear once,
    plus the filled KV rows at n_query_groups heads in bf16.  lm_head is replicated under TP (generate/tp.py:77-80)."""
    c = full_cfg
    per_layer = c.n_embd * (c.n_head + 2 * c.n_query_groups) * c.head_size + c.n_embd * c.n_embd + 3 * c.n_embd * c.intermediate_size
    head = c.padded_vocab_size * c.n_embd
    w = (c.n_layer * per_layer / tp + head / (tp if head_sharded else 1)) * BPP[mode]
    kv = 2 * c.n_layer * c.n_query_groups * c.head_size * 2 * ctx * batch / tp
    return {"weights": w, "kv": kv, "total": w + kv}


def gemv_only_step(model, x, y_attn, xn=None):
    """Exactly the dequant-GEMV launches of one fused decode step (norm prologues, residual / SwiGLU epilogues
    included), without attention / embedding / argmax: used to time the dominant kernel in isolation.
    xn: the pre-normalised rows for batches above model.SPLIT_NORM_ROWS, where the step runs RMSNorm as its own launch and the
    GEMVs take plain rows (GPT._fused_decode): the same launches as the step's, the norm launches themselves left out."""
    from lit_gpt_b200 import kernels as K

    eps = model.config.norm_eps

    def lin(q, inp, nxt=None, **kw):
        return q.gemv(inp, prefetch=None if nxt is None else nxt.tiled, **kw)

    def normed(q, inp, w, nxt=None, **kw):
        if xn is not None:
            return lin(q, xn, nxt, **kw)
        return lin(q, inp, nxt, norm_weight=w, norm_eps=eps, **kw)

    blocks = list(model.transformer.h)
    for li, block in enumerate(blocks):
        a, m = block.attn, block.mlp
        nxt_qkv = blocks[li + 1].attn.attn if li + 1 < len(blocks) else model.lm_head
        normed(a.attn, x, block.norm_1.weight, a.proj)
        if model.tp_world_size > 1:
            h = lin(a.proj, y_attn, m.fc_1)
            act = lin(m.fc_1, x, m.proj, other=m.fc_2, delta=h, norm_weight=block.norm_2.weight, norm_eps=eps)
            lin(m.proj, act, nxt_qkv)
        else:
            x2 = lin(a.proj, y_attn, m.fc_1, residual=x)
            act = normed(m.fc_1, x2, block.norm_2.weight, m.proj, other=m.fc_2)
            lin(m.proj, act, nxt_qkv, residual=x2)
    normed(model.lm_head, x, model.transformer.ln_f.weight, blocks[0].attn.attn)


def time_graph(graph, n: int, world: int) -> float:
    """ms for n replays, CUDA events on the launching stream, barrier + synchronize on both sides, max over ranks."""
    import torch.distributed as dist

    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        graph.replay()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.barrier()
    return float(ms.item())


def cpu_reference_sample(cfg_name: str, steps: int, layers: int = 4, warmup: int = 0) -> dict:
    """The reference's CPU path (generate/base.py on the host cores, fp32) restated by oracle/model_ref.py, on a
    BOUNDED sample: `layers` of the model's blocks at full width + the full lm_head, extrapolated to full depth."""
    from lit_gpt_b200 import Config
    from oracle import model_ref as M

    torch.set_num_threads(os.cpu_count() or 1)
    cfg = Config.from_name(cfg_name)
    rc = M.RefConfig.from_any(Config.from_name(cfg_name, n_layer=layers))
    g = torch.Generator().manual_seed(1234)
    p = {}
    shapes = {"attn.attn": ((cfg.n_head + 2 * cfg.n_query_groups) * cfg.head_size, cfg.n_embd), "attn.proj": (cfg.n_embd, cfg.n_embd),
              "mlp.fc_1": (cfg.intermediate_size, cfg.n_embd), "mlp.fc_2": (cfg.intermediate_size, cfg.n_embd),
              "mlp.proj": (cfg.n_embd, cfg.intermediate_size)}
    for i in range(layers):
        for k, s in shapes.items():
            p[f"transformer.h.{i}.{k}.weight"] = torch.empty(s).normal_(0, 0.02, generator=g)
        p[f"transformer.h.{i}.norm_1.weight"] = torch.ones(cfg.n_embd)
        p[f"transformer.h.{i}.norm_2.weight"] = torch.ones(cfg.n_embd)
    p["transformer.ln_f.weight"] = torch.ones(cfg.n_embd)
    p["transformer.wte.weight"] = torch.empty(cfg.padded_vocab_size, cfg.n_embd).normal_(0, 0.02, generator=g)
    p["lm_head.weight"] = torch.empty(cfg.padded_vocab_size, cfg.n_embd).normal_(0, 0.02, generator=g)
    T = 16
    st = M.RefState(rc, T + steps + warmup + 2, 1, torch.float32, pos_dtype=torch.float32)
    prompt = torch.randint(0, cfg.vocab_size, (T,), generator=g)
    with torch.inference_mode():
        logits = M.gpt_forward(rc, p, prompt.view(1, -1), torch.arange(T), st)
        tok = M.sample(logits, 0.0)
        # head-only time (ln_f + lm_head) to extrapolate depth correctly
        x = torch.randn(1, 1, cfg.n_embd)
        t0 = time.perf_counter()
        for _ in range(2):
            M.linear(M.rmsnorm(x, p["transformer.ln_f.weight"], cfg.norm_eps), p["lm_head.weight"])
        t_head = (time.perf_counter() - t0) / 2
        for i in range(warmup):
            logits = M.gpt_forward(rc, p, tok.view(1, 1), torch.tensor([T + i]), st)
            tok = M.sample(logits, 0.0)
        t0 = time.perf_counter()
        for i in range(steps):
            logits = M.gpt_forward(rc, p, tok.view(1, 1), torch.tensor([T + warmup + i]), st)
            tok = M.sample(logits, 0.0)
        t_step = (time.perf_counter() - t0) / steps
    full = layers >= cfg.n_layer
    t_full = t_step if full else (t_step - t_head) * cfg.n_layer / layers + t_head
    depth = "all blocks (full depth, nothing extrapolated)" if full else \
        f"{layers} of {cfg.n_layer} blocks at full width + full lm_head, extrapolated to full depth ({t_full * 1e3:.1f} ms/token)"
    return {"value": 1.0 / t_full, "unit": "tokens/s", "cores": torch.get_num_threads(), "kind": "port", "steps_timed": steps,
            "timed_s": t_step * steps,
            "sample": f"oracle/model_ref.py generate loop (reference generate/base.py restated), fp32, {depth}; "
                      f"{steps} decode tokens timed after a {T}-token prompt and {warmup} warm-up tokens; "
                      f"measured {t_step * 1e3:.1f} ms/token"}


def run_reference(args) -> None:
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    cfg_name, mode, batch, T, ctx = WORKLOADS[args.workload]
    from lit_gpt_b200 import Config

    # exactly `--steps` decode tokens are timed after `--warmup` untimed ones.  Full depth when the host has the memory for
    # the fp32 weights (7B: 27 GB) and the run stays within minutes; else a 4-block sample at full width, extrapolated
    # (stated in cpu_baseline.sample either way).
    cfg = Config.from_name(cfg_name)
    per_layer = cfg.n_embd * (cfg.n_head + 2 * cfg.n_query_groups) * cfg.head_size + cfg.n_embd * cfg.n_embd + 3 * cfg.n_embd * cfg.intermediate_size
    need = 4 * (cfg.n_layer * per_layer + 2 * cfg.padded_vocab_size * cfg.n_embd)
    try:
        import psutil

        avail = psutil.virtual_memory().available
    except Exception:
        avail = 0
    full_depth = avail > need * 1.3 and args.steps + args.warmup <= 64 and os.environ.get("B200_REF_SAMPLE", "0") != "1"
    steps = max(1, args.steps)
    t0 = time.perf_counter()
    cb = cpu_reference_sample(cfg_name, steps, layers=cfg.n_layer if full_depth else 4, warmup=max(0, args.warmup))
    line = {
        "impl": "reference", "metric": "Llama-2-7B nf4 decode tok/s (bs=1)", "value": cb["value"], "unit": "tokens/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": args.warmup, "ms_per_step": 1e3 / cb["value"],
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "note": "reference CPU path (bitsandbytes has no CPU nf4: fp32 weights)"},
        "cpu_baseline": cb, "e2e": {"value": cb["value"], "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "wall_s": time.perf_counter() - t0,
    }
    print(json.dumps(line), flush=True)


def run_ours(args) -> None:
    import torch.distributed as dist

    import lit_gpt_b200.generate.base as gb
    from lit_gpt_b200 import QuantPrecision
    from lit_gpt_b200 import kernels as K
    from lit_gpt_b200.fabric import Fabric

    if not torch.cuda.is_available():
        raise SystemExit("bench.py (impl ours) needs a CUDA device: the product path has no CPU fallback")
    if os.environ.get("B200_TUNE_OCC"):  # development: A/B switches of b200_qgemv_tiled_tune (e.g. 14 = no K / V cache L2 hint)
        K.qgemv_tiled_tune(0, int(os.environ["B200_TUNE_OCC"]), False)
    cfg_name, mode, batch, T, ctx = WORKLOADS[args.workload]
    world = int(os.environ.get("WORLD_SIZE", 1))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    fabric = Fabric(devices="auto", strategy="ddp" if world > 1 else None, plugins=QuantPrecision(mode, torch.bfloat16))
    fabric.launch()
    rank = fabric.global_rank
    K_steps, W = args.steps, max(args.warmup, 3)
    max_seq = ctx + T + W + K_steps + 2
    t_build = time.perf_counter()
    model, full_cfg = build_random_model(cfg_name, mode, max_seq, batch, fabric, n_layer=args.layers)
    torch.cuda.synchronize()
    t_build = time.perf_counter() - t_build
    assert model.fused_decode_ready(), "model is not on the fused CUDA path"
    dev = fabric.device
    g = torch.Generator().manual_seed(1234)
    prompt_host = torch.randint(0, full_cfg.vocab_size, (T,), generator=g).pin_memory()

    def prefill_context():
        for blk in model.transformer.h:
            blk.attn.kv_cache.reset_parameters()
        if ctx:  # synthetic N(0,1) K/V history, then the prompt after it
            gg = torch.Generator(device=dev).manual_seed(7)
            for blk in model.transformer.h:
                blk.attn.kv_cache.k[:, :, :ctx].normal_(generator=gg)
                blk.attn.kv_cache.v[:, :, :ctx].normal_(generator=gg)

    # ---- device-resident timing: K replays of the captured decode step --------------------------------------
    prefill_context()
    prompt = prompt_host.to(dev, non_blocking=True)
    tok = gb.next_token(model, torch.arange(ctx, ctx + T, device=dev), prompt.view(1, -1).expand(batch, -1).contiguous(), temperature=0.0)
    dg = gb.DecodeGraph(model, 0.0, None, torch.int64)
    if batch > 1:
        dg.token = torch.zeros(batch, dtype=torch.int64, device=dev)
        dg._step = lambda: K.argmax_(model(dg.token.view(batch, 1), dg.pos), dg.token, pos_inout=dg.pos)
    dg.token.copy_(tok.expand(batch) if batch > 1 else tok)
    dg.pos.fill_(ctx + T)
    n0 = K.launch_count()
    dg._step()
    launches_per_step = K.launch_count() - n0
    dg.pos.fill_(ctx + T)
    dg.capture()
    sampler = ClockSampler(dev)
    if rank == 0:
        sampler.start()   # running through the warm-up so that the first timed step is already covered
    time_graph(dg.graph, W, world)
    sampler.mark_start()
    ms = time_graph(dg.graph, K_steps, world)
    sampler.mark_end()
    clocks = sampler.stop() if rank == 0 else {}
    tok_s = batch * K_steps / (ms / 1e3)
    ctx_mid = ctx + T + W + K_steps // 2
    bytes_step = step_bytes(full_cfg, mode, ctx_mid, batch, world, bool(getattr(model, "lm_head_sharded", False)))

    # ---- dominant kernel in isolation: the dequant-GEMV launches of one step, same weights (3.7 GB >> L2) ----
    x = torch.randn(batch, model.transformer.wte.weight.shape[1], device=dev).to(torch.bfloat16)
    y_attn = torch.randn(batch, model.transformer.h[0].attn.proj.in_features, device=dev).to(torch.bfloat16)
    from lit_gpt_b200.model import SPLIT_NORM_ROWS

    split_norm = batch > SPLIT_NORM_ROWS and world == 1   # the step then normalises in its own launch and the GEMVs take plain rows
    xn = K.rmsnorm(x, model.transformer.ln_f.weight, model.config.norm_eps) if split_norm else None
    n0 = K.launch_count()
    gemv_only_step(model, x, y_attn, xn)
    gemv_launches = K.launch_count() - n0
    gg = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gg):
        gemv_only_step(model, x, y_attn, xn)
    time_graph(gg, W, 1)
    reps = max(8, min(K_steps, 64))
    gemv_ms = time_graph(gg, reps, 1) / reps
    peaks, peak_src = measured_peaks()
    achieved = bytes_step["weights"] / (gemv_ms / 1e3) / 1e9
    # dram__bytes_read + dram__bytes_write per launch from the committed ncu capture of this workload (ratio to the
    # algorithmic bytes, measured per layer shape); null for workloads that were not captured
    traffic = None
    tpath = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r02_traffic.json")
    if not os.path.exists(tpath):
        tpath = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r01_traffic.json")
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
        if tj.get("workload") == args.workload and world == 1:
            traffic = tj["traffic_over_algorithmic"] * bytes_step["weights"] / max(1, gemv_launches)
    kernel_name = "qgemv_v4_kernel" if batch == 1 or world > 1 else ("qgemv_nb_kernel" if split_norm else "qgemv_v4_kernel (norm-fused launches) + qgemv_nb_kernel (plain rows)")
    roofline = {"bound": "hbm", "kernel": kernel_name, "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                "frac": achieved / peaks["hbm_gbs"], "traffic": traffic,
                "traffic_source": None if traffic is None else "constant: ncu dram__bytes_read+write of the committed capture (profiles/), "
                                                                "scaled by this run's algorithmic bytes; not measured by this run",
                "peak_source": peak_src, "streamed_bytes_per_param": BPP_STREAMED[mode],
                "launches_per_step": gemv_launches, "avg_launch_us": gemv_ms * 1e3 / max(1, gemv_launches),
                "algorithmic_bytes_per_launch": bytes_step["weights"] / max(1, gemv_launches),
                "whole_step_frac": bytes_step["total"] / (ms / K_steps / 1e3) / 1e9 / peaks["hbm_gbs"],
                "frac_of_nominal_8TBs": achieved / 8000.0}

    # ---- end to end through generate(): pinned host prompt -> H2D -> prefill + K tokens -> D2H ------------------
    e2e = None
    if batch == 1 and ctx == 0:  # generate() starts at position 0: long-context points report the device number only
        prefill_context()
        n_total = T + K_steps
        gb.generate(model, prompt_host.to(dev), min(n_total, T + 4), temperature=0.0, eos_id=-1)  # warm
        prefill_context()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        out = gb.generate(model, prompt_host.to(dev, non_blocking=True), n_total, temperature=0.0, eos_id=-1)
        out_host = out.cpu()
        e1.record()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        e_ms = torch.tensor([max(e0.elapsed_time(e1), wall * 1e3)], device=dev)
        if world > 1:
            dist.all_reduce(e_ms, op=dist.ReduceOp.MAX)
        new = out_host.numel() - T
        polls = -(-max(new - 1, 1) // gb.EOS_POLL)
        e2e = {"value": new / (float(e_ms.item()) / 1e3), "unit": "tokens/s", "h2d_bytes_per_step": T * 8 / new,
               "d2h_bytes_per_step": (out_host.numel() * 8 + polls * (new // 2) * 8) / new,
               "note": f"generate(): {T}-token prompt from pinned host memory, prefill + {new} greedy tokens, EOS polled on the "
                       f"host every {gb.EOS_POLL} tokens, sequence copied back; includes python + launch overhead"}

    if rank != 0:
        return
    cb = None
    if world == 1 and not args.no_cpu_baseline:
        cb = cpu_reference_sample(cfg_name, 4)
    line = {
        "metric": "Llama-2-7B nf4 decode tok/s (bs=1)" if args.workload == "llama2-7b-nf4-bs1" else f"{args.workload} decode tok/s",
        "value": tok_s, "unit": "tokens/s", "n_gpus": world, "steps": K_steps, "warmup": W, "ms_per_step": ms / K_steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": args.workload, "model": cfg_name, "quantize": mode, "batch": batch, "prompt_tokens": T,
                   "context_prefill": ctx, "max_seq": max_seq, "parallelism": f"tp{world}", "lm_head_sharded": bool(getattr(model, "lm_head_sharded", False)), "n_layer": full_cfg.n_layer,
                   "l2_policy": "weights per step (>= 0.4 GB/GPU) exceed the 126 MB L2; no flush needed",
                   "bytes_per_step": bytes_step, "build_s": round(t_build, 1)},
        "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches_per_step * K_steps), "launches_per_step": int(launches_per_step),
        "roofline": roofline, "cpu_baseline": cb,
        "hbm_frac_of_measured_whole_step": roofline["whole_step_frac"],
    }
    print(json.dumps(line), flush=True)


def main() -> None:
    # stdout carries exactly one JSON line: NCCL's own banner / debug lines (when the box sets NCCL_DEBUG) go to stderr
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":  # the version banner is printf'd to stdout whatever the file is
        os.environ["NCCL_DEBUG"] = "WARN"
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=256)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default="llama2-7b-nf4-bs1")
    ap.add_argument("--layers", type=int, default=None, help="debug only: build fewer blocks (invalidates the number)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized():
        teardown()


def teardown() -> None:
    """Orderly multi-rank exit: captured graphs and peer mappings are dropped BEFORE the process group (a CUDA graph that
    captured an NCCL collective keeps the communicator busy: destroying the group first is what used to hang).  A watchdog
    turns a teardown that still does not finish within 30 s into a plain exit, after everything has been printed."""
    import gc

    import torch.distributed as dist

    sys.stdout.flush()
    sys.stderr.flush()

    def _bail():
        sys.stderr.write("bench.py: teardown did not finish within 30 s; leaving without destructors\n")
        sys.stderr.flush()
        os._exit(0)

    wd = threading.Timer(30.0, _bail)
    wd.daemon = True
    wd.start()
    torch.cuda.synchronize()
    dist.barrier()
    for obj in gc.get_objects():
        try:
            if type(obj).__name__ == "GPT" and hasattr(obj, "__dict__"):
                obj.__dict__.pop("_decode_graphs", None)
                comm = obj.__dict__.pop("_p2p", None)
                if comm:
                    comm.close()
        except Exception:
            pass
    gc.collect()
    torch.cuda.synchronize()
    dist.barrier()
    dist.destroy_process_group()
    wd.cancel()


if __name__ == "__main__":
    main()
