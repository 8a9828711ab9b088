"""Layer-wise placement of one model over the GPUs of a box -- ``generate/sequentially.py`` of the reference (lines 31-120).

Same contract (pinned by the reference's tests/test_generate_sequentially.py): ``layer_to_device`` maps block i to device
``i // (n_layer / devices)``, ``sequential`` materialises every block (and its kv cache) on its device and everything else on the
root, the only supported partitioning is the balanced one (same ``NotImplementedError``).  Weights are quantized by this library's
kernels ON the device a block lands on (``QLinear`` quantizes at the device move, like ``Params4bit``).

B200-native differences: the reference moves the activations with forward hooks around its un-fused modules; here the fused decode
chain (`GPT._fused_decode`: four GEMV launches + one attention launch per block) and the fused prefill run per device SEGMENT, and
the residual stream (8 KB at decode) hops between GPUs over NVLink with an asynchronous peer copy; each device holds its own copy
of the RoPE tables.  The hooks are still installed (same names, same places) so that the un-fused module path -- and the
reference's hook-inspection test -- behave as in the reference.  One process drives all devices, so the per-token loop is not a
single CUDA graph (`generate()` falls back to its eager loop for a model that spans devices).
"""
from __future__ import annotations

import itertools
import sys
import time
from collections import OrderedDict
from functools import partial
from pathlib import Path
from typing import Literal, Optional, Type

import torch

import lit_gpt_b200.generate.base as generate_base
from lit_gpt_b200 import GPT, Config, QuantPrecision
from lit_gpt_b200.fabric import Fabric, seed_everything
from lit_gpt_b200.model import Block, build_mask_cache
from lit_gpt_b200.utils import check_valid_checkpoint_dir, get_default_supported_precision


def _materialize_meta_tensors(module: torch.nn.Module, device: torch.device) -> None:
    """What lightning.fabric.utilities.init._materialize_meta_tensors does: leftover meta parameters / buffers get (uninitialised)
    storage on `device`."""
    for sub in module.modules():
        if any(t.is_meta for t in itertools.chain(sub.parameters(recurse=False), sub.buffers(recurse=False))):
            sub.to_empty(device=device, recurse=False)


@torch.inference_mode()
def sequential(model: GPT, root: torch.device, max_seq_length: int, devices: int):
    if model.config.n_layer % devices:
        # TODO: support smarter partitioning schemes
        raise NotImplementedError(
            f"Only balanced partitioning is implemented: n_layer={model.config.n_layer}, devices {devices}"
        )
    layers_per_rank = model.config.n_layer // devices
    # dictates where each block should be instantiated
    mapping = layer_to_device(model, chunk_on=Block, chunk_size=layers_per_rank)

    # the rope cache first (the kv caches are sized from it); rebuilt on the root below
    with root:
        model.max_seq_length = max_seq_length
        model.cos, model.sin = model.rope_cache()
    # materialize each block on the appropriate device
    for path, target_index in mapping.items():
        submodule = model.get_submodule(path)
        target_device = torch.device(root.type, target_index) if root.type != "cpu" else root
        print(f"Moving {path!r} to {target_device}", file=sys.stderr)
        with (torch.cuda.device(target_device) if target_device.type == "cuda" else _null()):
            # submodules loaded by the checkpoint will be on CPU (if no quantization). move them (QLinear quantizes here)
            replace_device(submodule, replace=torch.device("cpu"), by=target_device)
            # in case the checkpoint was partial, materialize leftover metas
            _materialize_meta_tensors(submodule, target_device)
            # and build the kv cache
            submodule.attn.kv_cache = submodule.attn.build_kv_cache(1, max_seq_length, model.cos.size(-1), target_device,
                                                                    model.transformer.wte.weight.dtype if target_device.type == "cuda" else None)
    # rebuild odd ends
    with root:
        # the mask cache which cannot be created with `set_kv_cache` because that will set it for all layers
        model.mask_cache = build_mask_cache(max_seq_length)
    # and everything that is not a block in the root
    _materialize_meta_tensors(model, root)
    replace_device(model, replace=torch.device("cpu"), by=root)
    model._fused_plan = None
    model.__dict__.pop("_decode_graphs", None)
    model.__dict__.pop("_prefill_graphs", None)
    model.__dict__.pop("_seq_tables", None)
    model.__dict__.pop("_spans", None)

    if devices > 1:
        # install hooks to move layer inputs/output between devices
        for layer_num, (path, target_index) in enumerate(mapping.items()):
            submodule = model.get_submodule(path)
            if layer_num >= layers_per_rank:
                # we need to move the block input on the boundaries between devices
                # and also on every non-root device because the RoPE and mask cache is shared
                target_device = torch.device(root.type, target_index)
                submodule.register_forward_pre_hook(partial(move_block_input, target_device))
            if layer_num == model.config.n_layer - 1:
                submodule.register_forward_hook(partial(move_block_output, root))

    return model


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


def layer_to_device(
    module: torch.nn.Module, chunk_on: Type[torch.nn.Module], chunk_size: int
) -> "OrderedDict[str, int]":
    """Create a mapping from layer (block) to device."""
    # this assumes that the definition order is the same as the execution order
    hits = [name for name, submodule in module.named_modules() if isinstance(submodule, chunk_on)]
    return OrderedDict((name, i // chunk_size) for i, name in enumerate(hits))


def move_block_input(device: torch.device, module: torch.nn.Module, ins):
    """``forward_pre_hook`` to move a Block's input before forward."""
    # during inference, none of the inputs are None: x, cos, sin, mask, input_pos
    return tuple(t.to(device) for t in ins)


def move_block_output(device: torch.device, module: torch.nn.Module, ins, outs) -> torch.Tensor:
    """``forward_hook`` to move a Block's output after forward."""
    return outs.to(device)


def replace_device(module: torch.nn.Module, replace: torch.device, by: torch.device) -> torch.nn.Module:
    for name, submodule in module.named_modules():
        tensors = dict(
            itertools.chain(submodule.named_parameters(recurse=False), submodule.named_buffers(recurse=False))
        )
        if not tensors:
            continue
        devices = {t.device for t in tensors.values()}
        if len(devices) != 1:
            # since this is using `submodule.to`, different devices in the same submodule is a problem
            path_to_device = {f"{name}.{p}": t.device for p, t in tensors.items()}
            raise ValueError(f"Found multiple devices: {path_to_device}")
        if devices.pop() == replace:
            submodule.to(by)
    return module


@torch.inference_mode()
def main(
    prompt: str = "What food do llamas eat?",
    *,
    num_samples: int = 1,
    max_new_tokens: int = 50,
    top_k: Optional[int] = 200,
    temperature: float = 0.8,
    checkpoint_dir: Path = Path("checkpoints/mistralai/Mistral-7B-Instruct-v0.1"),
    quantize: Optional[Literal["bnb.nf4", "bnb.nf4-dq", "bnb.fp4", "bnb.fp4-dq", "gptq.int4"]] = None,
    precision: Optional[str] = None,
    compile: bool = False,
) -> None:
    """Generates text samples with the model's blocks placed sequentially over all visible GPUs (one process)."""
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
    fabric = Fabric(devices=1, precision=precision, accelerator="cuda", plugins=plugins)
    total_devices = torch.cuda.device_count()
    print(f"Using {total_devices} devices", file=sys.stderr)
    check_valid_checkpoint_dir(checkpoint_dir)
    config = Config.from_json(checkpoint_dir / "lit_config.json")
    checkpoint_path = checkpoint_dir / "lit_model.pth"
    tokenizer = Tokenizer(checkpoint_dir)
    encoded = tokenizer.encode(prompt, device=fabric.device)
    prompt_length = encoded.size(0)
    max_returned_tokens = prompt_length + max_new_tokens

    print(f"Loading model {str(checkpoint_path)!r} with {config.__dict__}", file=sys.stderr)
    t0 = time.perf_counter()
    # not `init_module`: the Linear layers would be quantized on cuda:0 at checkpoint load; load on the CPU, convert, then place
    with fabric.init_tensor(), torch.device("meta"):
        model = GPT(config)
    print(f"Time to instantiate model: {time.perf_counter() - t0:.02f} seconds.", file=sys.stderr)
    t0 = time.perf_counter()
    state_dict = torch.load(str(checkpoint_path), mmap=True, map_location="cpu", weights_only=True)
    model.load_state_dict(state_dict, assign=True)
    print(f"Time to load the model weights: {time.perf_counter() - t0:.02f} seconds.", file=sys.stderr)
    if plugins is not None:
        model = plugins.convert_module(model)
    t0 = time.perf_counter()
    model = sequential(model, fabric.device, max_returned_tokens, total_devices)
    print(f"Time to sequential-ize the model: {time.perf_counter() - t0:.02f} seconds.", file=sys.stderr)
    if compile:
        raise NotImplementedError

    seed_everything(1234)
    for i in range(num_samples):
        t0 = time.perf_counter()
        y = generate_base.generate(
            model, encoded, max_returned_tokens, temperature=temperature, top_k=top_k, eos_id=tokenizer.eos_id
        )
        t = time.perf_counter() - t0
        for block in model.transformer.h:
            block.attn.kv_cache.reset_parameters()
        print(tokenizer.decode(y))
        tokens_generated = y.size(0) - prompt_length
        print(
            f"Time for inference {i + 1}: {t:.02f} sec total, {tokens_generated / t:.02f} tokens/sec", file=sys.stderr
        )
    print(f"Memory used: {torch.cuda.max_memory_allocated() / 1e9:.02f} GB", file=sys.stderr)
