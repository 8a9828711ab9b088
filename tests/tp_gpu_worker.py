"""Launched by tests/test_gpu_model.py::test_tp2_matches_single_gpu under torchrun (2 ranks, NCCL over NVLink).

A quantized model sharded as generate/tp.py:48-92 prescribes (shard in high precision, then quantize each shard on its
GPU, tp.py:166-190) must give the logits and the greedy rollout of the same model on one GPU: column-parallel shards
keep whole rows and row-parallel shards cut K at a multiple of the quantisation block, so the dequantised weights are
identical and only the order of the partial sums differs.
"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, sys.argv[1])

from lit_gpt_b200 import GPT, Config, QuantPrecision  # noqa: E402
from lit_gpt_b200.fabric import Fabric  # noqa: E402
from lit_gpt_b200.generate.tp import tensor_parallel  # noqa: E402
import lit_gpt_b200.generate.base as gb  # noqa: E402


def build(fabric, cfg_kw, sd, shard):
    """The load flow of generate/tp.py:163-192: weights on the CPU in 16 bit -> swap in QLinear -> shard -> move
    (the move quantizes each shard on its GPU)."""
    prev = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    try:
        model = GPT(Config.from_name("Llama-2-7b-hf", **cfg_kw)).eval()
    finally:
        torch.set_default_dtype(prev)
    model.load_state_dict(sd)
    model = fabric._precision.convert_module(model)
    if shard:
        model = tensor_parallel(fabric, model)
    with fabric.init_tensor():
        model.max_seq_length = 48
        model.cos, model.sin = model.rope_cache()
        model.set_kv_cache(batch_size=1)
    return fabric.to_device(model)


def main() -> None:
    # third case: intermediate_size / 2 = 1056 is not a multiple of the 64-element block -> the mlp.proj shard has no tiled
    # layout (generic kernel, its sum by NCCL) while the attention branch keeps the peer-memory hand-off: the mixed path
    for mode, inter in (("nf4", 2048), ("int4", 2048), ("nf4", 2112)):
        fabric = Fabric(devices="auto", strategy="ddp", plugins=QuantPrecision(mode, torch.bfloat16))
        fabric.launch()
        cfg_kw = dict(n_layer=2, n_head=8, n_query_groups=4, n_embd=1024, intermediate_size=inter, padded_vocab_size=1024, block_size=64)
        torch.manual_seed(0)
        ref = GPT(Config.from_name("Llama-2-7b-hf", **cfg_kw))
        ref.apply(ref._init_weights)
        sd = {k: v.to(torch.bfloat16) for k, v in ref.state_dict().items()}
        del ref
        prompt = torch.tensor([5, 17, 3, 99, 42, 7], device=fabric.device)
        with torch.inference_mode():
            single = build(fabric, cfg_kw, sd, shard=False)
            y1 = single(prompt.view(1, -1), torch.arange(6, device=fabric.device)).float()
            d1 = single(torch.tensor([[11]], device=fabric.device), torch.tensor([6], device=fabric.device)).float()
            for blk in single.transformer.h:
                blk.attn.kv_cache.reset_parameters()
            t1 = gb.generate(single, prompt, 24, temperature=0.0)
            del single
            sharded = build(fabric, cfg_kw, sd, shard=True)
            assert sharded.tp_world_size == 2
            assert (sharded.transformer.h[0].mlp.proj.tiled is None) == (inter == 2112)
            y2 = sharded(prompt.view(1, -1), torch.arange(6, device=fabric.device)).float()
            d2 = sharded(torch.tensor([[11]], device=fabric.device), torch.tensor([6], device=fabric.device)).float()
            for blk in sharded.transformer.h:
                blk.attn.kv_cache.reset_parameters()
            t2 = gb.generate(sharded, prompt, 24, temperature=0.0)
            d3 = None
            if inter == 2112:  # the same sharded model with every sum by NCCL: the mixed hand-off must agree with it
                os.environ["B200_TP_NCCL"] = "1"
                sharded.__dict__.pop("_p2p", None)
                sharded.__dict__.pop("_decode_graphs", None)
                for blk in sharded.transformer.h:
                    blk.attn.kv_cache.reset_parameters()
                sharded(prompt.view(1, -1), torch.arange(6, device=fabric.device))
                d3 = sharded(torch.tensor([[11]], device=fabric.device), torch.tensor([6], device=fabric.device)).float()
                os.environ["B200_TP_NCCL"] = "0"
        scale = y1.abs().max().item()
        e_p, e_d = (y1 - y2).abs().max().item(), (d1 - d2).abs().max().item()
        tol = 2 ** -5 * scale   # bf16 partial sums re-ordered across two ranks
        if inter == 2112:
            # the row-parallel shard's 64-element blocks straddle rows (flat quantisation of the shard, as bitsandbytes does), so
            # its code words differ from the unsharded matrix: only a loose bound against the single GPU, a tight one against NCCL
            tol = 2 ** -2 * scale
            e_n = (d2 - d3).abs().max().item()
            assert e_n <= 2 ** -6 * scale, (mode, inter, e_n, scale)
        assert e_p <= tol and e_d <= tol, (mode, e_p, e_d, tol)
        # greedy tokens may only differ where the single-GPU top-2 margin is inside the tolerance; compare the prefix
        same = int((t1 == t2).int().cumprod(0).sum())
        assert same >= (8 if inter != 2112 else 1), (mode, t1.tolist(), t2.tolist())
        # all ranks hold identical logits (NCCL all-reduce is bitwise identical across ranks; no broadcast, tp.py:200)
        gathered = [torch.empty_like(d2) for _ in range(2)]
        dist.all_gather(gathered, d2)
        assert torch.equal(gathered[0], gathered[1])
        if fabric.global_rank == 0:
            print(f"[tp2/{mode}/I={inter}] prefill err {e_p:.4f} decode err {e_d:.4f} (tol {tol:.4f}, |logit| max {scale:.2f}); greedy prefix {same}/24", flush=True)
        del sharded
    torch.cuda.synchronize()
    dist.barrier()
    if int(os.environ.get("RANK", 0)) == 0:
        print("TP_GPU_OK", flush=True)
    sys.stdout.flush()
    os._exit(0)  # NCCL communicators captured in CUDA graphs make interpreter teardown hang; everything is flushed


if __name__ == "__main__":
    main()
