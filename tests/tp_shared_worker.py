"""Launched by tests/test_gpu_tp_shared.py under torchrun: WORLD ranks that all drive cuda:0 (B200_TP_SHARED_DEVICE=1, gloo
control plane), so that a ONE-GPU box exercises the tensor-parallel decode path end to end: shard-then-quantize
(generate/tp.py:28-92, 163-192), the peer-memory hand-off inside the GEMV kernels (b200_tp_handoff: CUDA-IPC slots,
self-validating packets, buffer alternation, tag wrap-around) and the captured decode graph.

The checker is the ORACLE, not this repo's single-GPU model: the reference model restated in oracle/model_ref.py, run
unsharded in fp32 with every Linear weight rebuilt from the per-rank shards exactly as each rank holds them,
W' = concat_r dequant(quant(shard_r(W))) (oracle/quant_ref: rows for column-parallel, columns for row-parallel Linears).
Tolerance = the rule of tests/test_gpu_model.py: 2 x (error of the bf16 reference formulation) + 1e-3 * max|logit|;
greedy tokens must agree on every step whose oracle top-1 / top-2 margin exceeds it.

argv: repo_root mode n_head n_query_groups [seq0 [intermediate_size]]
"""
import os
import sys

import torch
import torch.distributed as dist

ROOT = sys.argv[1]
sys.path.insert(0, ROOT)

from lit_gpt_b200 import GPT, Config, QuantPrecision  # noqa: E402
from lit_gpt_b200.fabric import Fabric  # noqa: E402
from lit_gpt_b200.generate.tp import shard_lm_head, tensor_parallel  # noqa: E402
import lit_gpt_b200.generate.base as gb  # noqa: E402
from oracle import model_ref as M  # noqa: E402
from oracle import quant_ref as Q  # noqa: E402
from oracle import quant_ref_c as QC  # noqa: E402

STYLE = {"attn.attn": "colwise", "attn.proj": "rowwise", "mlp.fc_1": "colwise", "mlp.fc_2": "colwise", "mlp.proj": "rowwise"}


def qdq(w: torch.Tensor, mode: str) -> torch.Tensor:
    w = w.to(torch.bfloat16).float().contiguous().numpy()
    if mode == "int4":
        qw, s, z = Q.quantize_int4_g128(w)
        return torch.from_numpy(QC.dequantize_int4(qw, s, z, w.shape[1]))
    pk, am = QC.quantize_4bit(w, mode)
    return torch.from_numpy(QC.dequantize_4bit(pk, am, w.shape, mode))


def oracle_tp_weights(p, mode, world):
    """What the WORLD ranks hold, reassembled: each shard quantized on its own (tp.py:163-192 shards first)."""
    out = {}
    for k, v in p.items():
        if v.dim() == 2 and "wte" not in k:
            style = next((s for n, s in STYLE.items() if k.endswith(n + ".weight")), None)
            if style is None:  # lm_head is replicated (tp.py:77-80)
                out[k] = qdq(v, mode)
            else:
                parts = [qdq(M.tensor_parallel_shard(v, world, r, style), mode) for r in range(world)]
                out[k] = torch.cat(parts, dim=0 if style == "colwise" else 1)
        else:
            out[k] = v.to(torch.bfloat16).float()
    return out


def raw_u32(ptr: int, n: int, device) -> torch.Tensor:
    class _P:
        __cuda_array_interface__ = {"shape": (n,), "typestr": "<i4", "data": (ptr, False), "version": 2}

    return torch.as_tensor(_P(), device=device)


def main() -> None:
    mode, n_head, n_groups = sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
    seq0 = int(sys.argv[5]) if len(sys.argv) > 5 else 0
    inter = int(sys.argv[6]) if len(sys.argv) > 6 else 0
    fabric = Fabric(devices="auto", strategy="ddp", plugins=QuantPrecision(mode, torch.bfloat16))
    fabric.launch()
    world, rank = fabric.world_size, fabric.global_rank
    shared = os.environ.get("B200_TP_SHARED_DEVICE", "0") == "1"
    if shared:
        assert fabric.device == torch.device("cuda", 0) and dist.get_backend() == "gloo"
    else:  # one GPU per rank, NCCL control plane (tests/test_gpu_model.py::test_tp2_matches_oracle_multi_gpu)
        assert fabric.device == torch.device("cuda", fabric.local_rank) and dist.get_backend() == "nccl"
    cfg_kw = dict(n_layer=2, n_head=n_head, n_query_groups=n_groups, n_embd=128 * n_head, intermediate_size=inter or 256 * world * 2,
                  padded_vocab_size=1024, vocab_size=1024, block_size=64)
    cfg = Config.from_name("Llama-2-7b-hf", **cfg_kw)
    rc = M.RefConfig.from_any(cfg)
    p = M.init_weights(rc, seed=4321, dtype=torch.bfloat16)
    n_total, T = 30, 6
    prompt = torch.tensor([5, 17, 3, 99, 42, 7])

    prev = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    try:
        model = GPT(Config.from_name("Llama-2-7b-hf", **cfg_kw)).eval()
    finally:
        torch.set_default_dtype(prev)
    model.load_state_dict(p)
    model = fabric._precision.convert_module(model)
    model = tensor_parallel(fabric, model)
    head_sharded = os.environ.get("B200_TP_SHARD_LM_HEAD", "0") == "1"
    if head_sharded:   # lm_head sharded by rows; the oracle is unchanged (whole rows per shard: same code words)
        model = shard_lm_head(fabric, model)
        assert model.lm_head.out_features == cfg.padded_vocab_size // world
    with fabric.init_tensor():
        model.max_seq_length = 48
        model.cos, model.sin = model.rope_cache()
        model.set_kv_cache(batch_size=1)
    model = fabric.to_device(model)
    assert model.tp_world_size == world and model.config.n_query_groups == n_groups // world
    if inter and (inter // world) % 64:
        assert model.transformer.h[0].mlp.proj.fold == 2 and model.transformer.h[0].mlp.proj.tiled is not None

    with torch.inference_mode():
        dev = fabric.device
        # ---- oracle (every rank computes it: identical inputs, no communication) ------------------------------------
        pw = oracle_tp_weights(p, mode, world)
        cos, sin = model.cos.cpu(), model.sin.cpu()

        def run_oracle(dtype, c, s, forced=None):
            st = M.RefState(rc, n_total, 1, dtype, cos=c, sin=s)
            toks, logits = M.generate(rc, {k: v.to(dtype) for k, v in pw.items()}, st, prompt, n_total, temperature=0.0,
                                      return_logits=True, forced=forced)
            return toks, logits.float()

        toks32, logits32 = run_oracle(torch.float32, cos.float(), sin.float())
        _, logits16 = run_oracle(torch.bfloat16, cos, sin, forced=toks32)  # same tokens: a rounding error, not a diverged rollout
        # ---- sharded CUDA path, teacher-forced along the oracle's tokens ---------------------------------------------
        got = [model(prompt.view(1, -1).to(dev), torch.arange(T, device=dev))[0, -1].float().cpu()]
        first = model(toks32[T].view(1, 1).to(dev), torch.tensor([T], device=dev))  # creates the peer communicator
        comm = model.__dict__.get("_p2p")
        assert comm, "the fused decode step must run on the peer-memory hand-off, not on the all-reduce path"
        if seq0:  # start both buffers' sequence counters just below the 16-bit tag wrap (tag = seq % 65535 + 1)
            torch.cuda.synchronize()
            dist.barrier()
            for b in (0, 1):
                raw_u32(comm._flags(comm._local, b), 1, dev).fill_(seq0)
            torch.cuda.synchronize()
            dist.barrier()
            first = model(toks32[T].view(1, 1).to(dev), torch.tensor([T], device=dev))
        got.append(first[0, -1].float().cpu())
        for i in range(1, n_total - T - 1):
            got.append(model(toks32[T + i].view(1, 1).to(dev), torch.tensor([T + i], device=dev))[0, -1].float().cpu())
        got = torch.stack(got)
        scale = logits32.abs().max().item()
        e_ref = (logits16 - logits32).abs().max().item()
        e_k = (got - logits32).abs().max().item()
        tol = 2.0 * e_ref + 1e-3 * scale
        assert e_k <= tol, (mode, world, e_k, tol)
        top2 = logits32.topk(2, dim=-1).values
        decisive = (top2[:, 0] - top2[:, 1]) > tol
        agree = got.argmax(-1) == logits32.argmax(-1)
        assert bool(agree[decisive].all()), agree.tolist()
        # ---- every rank holds the same logits bit for bit (no broadcast in tp.py:200) ----------------------------------
        got_d = got if shared else got.to(dev)   # gloo gathers host tensors, NCCL device tensors
        gathered = [torch.empty_like(got_d) for _ in range(world)]
        dist.all_gather(gathered, got_d)
        assert all(torch.equal(gathered[0], t) for t in gathered)
        # ---- free-running generate(): the captured decode graph with the hand-offs inside ---------------------------
        for blk in model.transformer.h:
            blk.attn.kv_cache.reset_parameters()
        out = gb.generate(model, prompt.to(dev), n_total, temperature=0.0).cpu()
        assert torch.equal(out[:T], prompt)
        div = next((i for i in range(T, n_total) if out[i].item() != toks32[i].item()), None)
        if div is not None:
            assert not bool(decisive[div - T]), f"rollout diverged at decisive step {div}"
        out_d = out if shared else out.to(dev)
        outs = [torch.empty_like(out_d) for _ in range(world)]
        dist.all_gather(outs, out_d)
        assert all(torch.equal(outs[0], t) for t in outs)
        e_n = None
        if not shared:
            # the same sharded model with every row-parallel sum by NCCL all-reduce (generate/tp.py:73-74): the peer-memory
            # hand-off must agree with it to bf16 rounding of the partial sums
            os.environ["B200_TP_NCCL"] = "1"
            model.__dict__.pop("_p2p", None)
            model.__dict__.pop("_decode_graphs", None)
            for blk in model.transformer.h:
                blk.attn.kv_cache.reset_parameters()
            model(prompt.view(1, -1).to(dev), torch.arange(T, device=dev))
            d_nccl = model(toks32[T].view(1, 1).to(dev), torch.tensor([T], device=dev))[0, -1].float().cpu()
            assert model.__dict__.get("_p2p") is False
            os.environ["B200_TP_NCCL"] = "0"
            e_n = (d_nccl - got[1]).abs().max().item()
            assert e_n <= 2 ** -6 * scale, (e_n, scale)
        seq_now = int(raw_u32(comm._flags(comm._local, 0), 1, dev).item())
    if rank == 0:
        print(f"[tp{world}/{'shared-gpu' if shared else 'multi-gpu'}/{mode}/H{n_head}G{n_groups}/I{cfg.intermediate_size}{'/lm_head-sharded' if head_sharded else ''}] bf16-reference err {e_ref:.4f} cuda err {e_k:.4f} tol {tol:.4f} "
              f"(|logit| max {scale:.2f}); top-1 {int(agree.sum())}/{len(agree)} (decisive {int(decisive.sum())}); "
              f"greedy prefix {(div or n_total) - T}/{n_total - T}; hand-off seq {seq0} -> {seq_now}"
              + ("" if e_n is None else f"; |p2p - nccl| {e_n:.4f}"), flush=True)
    if seq0:
        assert seq_now > seq0 and (seq0 % 65535) > (seq_now % 65535), "the run was meant to cross the tag wrap-around"
    torch.cuda.synchronize()
    dist.barrier()
    model.__dict__.pop("_decode_graphs", None)
    comm.close()
    if rank == 0:
        print("TP_SHARED_OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
