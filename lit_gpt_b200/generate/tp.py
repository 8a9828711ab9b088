"""Tensor parallelism across the GPUs of one NVSwitch box — ``generate/tp.py`` of the reference (lines 28-215).

Same sharding contract (pinned by the reference's tests/test_generate_tp.py): ``attn.attn`` / ``mlp.fc_1`` /
``mlp.fc_2`` column-parallel, both ``proj`` row-parallel, ``config.n_head / n_embd / n_query_groups`` divided by the
world size, weights sharded in full precision FIRST and quantized per rank when moved to the GPU (tp.py:163-192).
One exchange per row-parallel output: a sum all-reduce over NCCL / NVLink.  On the fused decode path the all-reduce is
issued between the row-parallel GEMV and the next GEMV, whose prologue adds the reduced branch to the residual stream,
so the whole step (collectives included) is one CUDA graph.
"""
from __future__ import annotations

import sys
import time
from functools import partial
from pathlib import Path
from typing import Any, Literal, Optional, Union

import torch
import torch.distributed as dist

import lit_gpt_b200.generate.base as generate_base
from lit_gpt_b200 import GPT, Config, QuantPrecision
from lit_gpt_b200.fabric import Fabric, seed_everything
from lit_gpt_b200.model import CausalSelfAttention, GptNeoxMLP, LLaMAMLP
from lit_gpt_b200.utils import check_valid_checkpoint_dir, get_default_supported_precision


def tensor_parallel_linear(fabric: Any, linear: torch.nn.Module, style: str) -> None:
    world_size = fabric.world_size
    dim, attr = {"colwise": (0, "out_features"), "rowwise": (1, "in_features")}[style]
    size = getattr(linear, attr)
    if size % world_size != 0:
        raise ValueError(
            f"This linear's {attr} value ({size}) is not evenly divisible by the world size ({world_size})"
        )
    if getattr(linear, "is_quantized", False):
        raise RuntimeError("shard before quantizing: quantization blocks follow the SHARD's flattened order")
    shard = torch.tensor_split(linear.weight, world_size, dim=dim)[fabric.global_rank]
    linear.weight.data = shard  # `.data` assignment keeps the (Q)Linear's parameter object
    setattr(linear, attr, shard.size(dim))
    if linear.bias is not None and dim == 0:
        shard = torch.tensor_split(linear.bias, world_size)[fabric.global_rank]
        linear.bias = torch.nn.Parameter(shard, requires_grad=linear.bias.requires_grad)


def tensor_parallel_mlp(fabric: Any, mlp: Union[GptNeoxMLP, LLaMAMLP]) -> None:
    if isinstance(mlp, LLaMAMLP):
        tensor_parallel_linear(fabric, mlp.fc_1, "colwise")
        tensor_parallel_linear(fabric, mlp.fc_2, "colwise")
        tensor_parallel_linear(fabric, mlp.proj, "rowwise")
        mlp.register_forward_hook(partial(all_reduce_output, fabric.world_size))
    elif isinstance(mlp, GptNeoxMLP):
        tensor_parallel_linear(fabric, mlp.fc, "colwise")
        tensor_parallel_linear(fabric, mlp.proj, "rowwise")
        mlp.register_forward_hook(partial(all_reduce_output, fabric.world_size))
    else:
        raise NotImplementedError


def tensor_parallel_attn(fabric: Any, attn: CausalSelfAttention) -> None:
    tensor_parallel_linear(fabric, attn.attn, "colwise")
    tensor_parallel_linear(fabric, attn.proj, "rowwise")
    attn.register_forward_hook(partial(all_reduce_output, fabric.world_size))


def all_reduce_output(world_size: int, module: torch.nn.Module, ins: Any, outs: torch.Tensor) -> torch.Tensor:
    """Module-path hook (prefill, non-fused models): sum the row-parallel partial outputs over all ranks."""
    if world_size > 1 and dist.is_available() and dist.is_initialized():
        outs = outs.contiguous()
        dist.all_reduce(outs, op=dist.ReduceOp.SUM)
    return outs


def all_gather_output(world_size: int, module: torch.nn.Module, ins: Any, outs: torch.Tensor) -> torch.Tensor:
    """Module-path hook of a row-sharded lm_head (prefill, non-fused models): concatenate the ranks' vocabulary slices."""
    if world_size > 1 and dist.is_available() and dist.is_initialized():
        outs = outs.contiguous()
        parts = [torch.empty_like(outs) for _ in range(world_size)]
        dist.all_gather(parts, outs)
        outs = torch.cat(parts, dim=-1)
    return outs


def shard_lm_head(fabric: Any, model: GPT) -> GPT:
    """NOT in the reference (generate/tp.py:77-80 leaves `lm_head` replicated; SURVEY §8f-3 lists this as the next step): shard
    `lm_head` by rows (vocabulary slices) so that each rank streams 1 / world of it.  Call after `tensor_parallel`, before the
    move to the GPU.  Decode: the slices are all-gathered over the peer-memory slots (b200_tp_gather) and every rank holds the full
    logits, bit-identical to the replicated Linear; prefill / module path: `all_gather` hook.  Skipped (model unchanged) when the
    vocabulary does not split into 16-row tiles."""
    world_size = fabric.world_size
    head = model.lm_head
    if world_size < 2 or head.out_features % world_size or (head.out_features // world_size) % 16 or head.bias is not None:
        return model
    tensor_parallel_linear(fabric, head, "colwise")
    head.register_forward_hook(partial(all_gather_output, world_size))
    model.lm_head_sharded = True
    model._fused_plan = None
    return model


def tensor_parallel(fabric: Any, model: GPT) -> GPT:
    for block in model.transformer.h:
        tensor_parallel_mlp(fabric, block.mlp)
        tensor_parallel_attn(fabric, block.attn)

    world_size = fabric.world_size
    for attr in ("n_head", "n_embd", "n_query_groups"):
        size = getattr(model.config, attr)
        if size % world_size != 0:
            raise ValueError(f"This {attr} value ({size}) is not evenly divisible by the world size ({world_size})")
        setattr(model.config, attr, size // world_size)
    model.tp_world_size = world_size
    model.tp_rank = fabric.global_rank  # recorded in (and checked against) pre-quantized shard checkpoints
    model.tp_group = None  # default process group
    model._fused_plan = None
    return model


@torch.inference_mode()
def main(
    prompt: str = "What food do llamas eat?",
    *,
    num_samples: int = 1,
    max_new_tokens: int = 50,
    top_k: Optional[int] = 200,
    temperature: float = 0.8,
    checkpoint_dir: Path = Path("checkpoints/stabilityai/stablelm-base-alpha-3b"),
    quantize: Optional[Literal["bnb.nf4", "bnb.nf4-dq", "bnb.fp4", "bnb.fp4-dq", "gptq.int4"]] = None,
    precision: Optional[str] = None,
    compile: bool = False,
) -> None:
    """Generates text samples with the model sharded over all visible GPUs (launch with torchrun)."""
    from lit_gpt_b200.tokenizer import Tokenizer

    precision = precision or get_default_supported_precision(training=False)

    plugins = None
    if quantize is not None:
        if compile:
            raise NotImplementedError  # untested
        if "mixed" in precision:
            raise ValueError("Quantization and mixed precision is not supported.")
        dtype = {"16-true": torch.float16, "bf16-true": torch.bfloat16, "32-true": torch.float32}[precision]
        plugins = QuantPrecision(quantize, dtype)
        precision = None

    fabric = Fabric(devices="auto", strategy="ddp", precision=precision, plugins=plugins)
    fabric.launch()

    check_valid_checkpoint_dir(checkpoint_dir)
    config = Config.from_json(checkpoint_dir / "lit_config.json")
    checkpoint_path = checkpoint_dir / "lit_model.pth"

    tokenizer = Tokenizer(checkpoint_dir)
    encoded = tokenizer.encode(prompt, device=fabric.device)
    prompt_length = encoded.size(0)
    max_returned_tokens = prompt_length + max_new_tokens

    fabric.print(f"Loading model {str(checkpoint_path)!r} with {config.__dict__}", file=sys.stderr)
    t0 = time.perf_counter()
    with fabric.init_tensor(), torch.device("meta"):
        model = GPT(config)
    fabric.print(f"Time to instantiate model: {time.perf_counter() - t0:.02f} seconds.", file=sys.stderr)

    # ranks take turns: load on CPU -> convert -> shard -> move (quantizes the shard on its GPU); bounds host RAM
    for rank in range(fabric.world_size):
        if fabric.global_rank == rank:
            t0 = time.perf_counter()
            state_dict = torch.load(str(checkpoint_path), mmap=True, map_location="cpu", weights_only=True)
            model.load_state_dict(state_dict, assign=True)
            print(f"[{rank}] Time to load the model weights: {time.perf_counter() - t0:.02f} seconds.", file=sys.stderr)

            if fabric._precision is not None:
                model = fabric._precision.convert_module(model)

            t0 = time.perf_counter()
            model = tensor_parallel(fabric, model)
            import os

            if fabric._precision is not None and os.environ.get("B200_TP_SHARD_LM_HEAD", "1") == "1":
                model = shard_lm_head(fabric, model)  # beyond the reference (tp.py:77-80 replicates it): vocabulary slices per rank
            print(f"[{rank}] Time to tensor-parallelize the model: {time.perf_counter() - t0:.02f} seconds.", file=sys.stderr)

            with fabric.init_tensor():
                model.max_seq_length = max_returned_tokens
                model.cos, model.sin = model.rope_cache()  # the tables were created on the meta device
                model.set_kv_cache(batch_size=1)
            model.eval()

            t0 = time.perf_counter()
            model = fabric.to_device(model)
            print(f"[{rank}] Time to move the model: {time.perf_counter() - t0:.02f} seconds.", file=sys.stderr)
        fabric.barrier()

    seed_everything(1234)
    for i in range(num_samples):
        t0 = time.perf_counter()
        y = generate_base.generate(
            model, encoded, max_returned_tokens, temperature=temperature, top_k=top_k, eos_id=tokenizer.eos_id
        )
        t = time.perf_counter() - t0
        for block in model.transformer.h:
            block.attn.kv_cache.reset_parameters()
        fabric.print(tokenizer.decode(y))
        tokens_generated = y.size(0) - prompt_length
        fabric.print(
            f"Time for inference {i + 1}: {t:.02f} sec total, {tokens_generated / t:.02f} tokens/sec", file=sys.stderr
        )
    if fabric.device.type == "cuda":
        fabric.print(f"Memory used: {torch.cuda.max_memory_allocated() / 1e9:.02f} GB", file=sys.stderr)


if __name__ == "__main__":
    torch.set_float32_matmul_precision("high")
    generate_base._cli(main)
