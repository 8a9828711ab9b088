"""Pre-quantized checkpoints.

The reference re-quantizes from 16-bit weights on every launch and, under tensor parallelism, serialises the ranks while
doing so (generate/tp.py:163-192; lit_gpt/utils.py:350-356 only knows plain ``weight`` / ``bias`` keys).  Here a model
that has been quantized once is written out as it sits in HBM -- packed codes plus statistics per ``QLinear`` in the
reference-layout form (bitsandbytes 0.41 nf4/fp4 bytes + absmax, the nested 8-bit statistics for ``-dq``, int8 ``CB`` +
``SCB``, or this project's int4 ``qweight / scales / qzeros``), everything else (embedding, norms) as plain tensors --
and read back without touching the quantizer: only the tiled kernel layout, a byte permutation, is rebuilt on load.
Under TP each rank saves and loads its own shard (``rank`` / ``world`` are recorded and checked).
"""
from __future__ import annotations

import os
from dataclasses import asdict
from typing import Any, Dict, Union

import torch
import torch.nn as nn

from ._cabi import B200Error
from .quant import QLinear

FORMAT = "b200gpt-quantized-v1"


def quantized_state(model: nn.Module) -> Dict[str, Any]:
    packed, plain = {}, {}
    qnames = []
    for name, mod in model.named_modules():
        if isinstance(mod, QLinear):
            packed[name] = mod.export_packed()
            qnames.append(name + ".")
    for k, v in model.state_dict().items():
        if not any(k.startswith(q) for q in qnames) and "kv_cache" not in k:
            plain[k] = v.detach().cpu()
    cfg = getattr(model, "config", None)
    return {"format": FORMAT, "config": asdict(cfg) if cfg is not None else None, "packed": packed, "plain": plain,
            "tp": {"world": int(getattr(model, "tp_world_size", 1)), "rank": int(getattr(model, "tp_rank", 0))}}


def save_quantized(model: nn.Module, path: Union[str, os.PathLike]) -> None:
    torch.save(quantized_state(model), str(path))


def load_quantized(model: nn.Module, path: Union[str, os.PathLike], device: Union[str, torch.device] = "cuda") -> nn.Module:
    """``model``: a GPT whose Linears are ``QLinear`` of the checkpoint's mode (``fabric.init_module`` / ``convert_module``), on
    meta, CPU or CUDA, already sharded if the checkpoint is a TP shard.  Returns the model on ``device`` ready to decode."""
    st = torch.load(str(path), map_location="cpu", weights_only=True)  # tensors / str / int / float / list / dict only
    if not isinstance(st, dict) or st.get("format") != FORMAT:
        raise B200Error(f"{path}: not a {FORMAT} checkpoint")
    world = int(getattr(model, "tp_world_size", 1))
    if st["tp"]["world"] != world:
        raise B200Error(f"checkpoint was written for tensor-parallel world size {st['tp']['world']}, the model is set up for {world}")
    rank = int(getattr(model, "tp_rank", 0))
    if st["tp"]["rank"] != rank:
        raise B200Error(f"checkpoint holds the shard of tensor-parallel rank {st['tp']['rank']}, this process is rank {rank}")
    dev = torch.device(device)
    mods = {n: m for n, m in model.named_modules() if isinstance(m, QLinear)}
    missing = sorted(set(mods) - set(st["packed"]))
    unexpected = sorted(set(st["packed"]) - set(mods))
    if missing or unexpected:
        raise B200Error(f"quantized modules do not match the checkpoint: missing {missing[:4]}, unexpected {unexpected[:4]}")
    for name, mod in mods.items():
        mod.import_packed(st["packed"][name], dev)
    # everything else: materialise on the device and load by name
    own = dict(model.named_parameters(recurse=True))
    own.update(dict(model.named_buffers(recurse=True)))
    qprefix = tuple(n + "." for n in mods)
    for k, v in st["plain"].items():
        if k.startswith(qprefix):
            continue
        parent, _, leaf = k.rpartition(".")
        m = model.get_submodule(parent) if parent else model
        cur = getattr(m, leaf, None)
        t = v.to(dev)
        if isinstance(cur, nn.Parameter) or leaf in getattr(m, "_parameters", {}):
            setattr(m, leaf, nn.Parameter(t, requires_grad=False))
        elif leaf in getattr(m, "_buffers", {}):
            m._buffers[leaf] = t
        else:
            raise B200Error(f"unexpected tensor {k!r} in the checkpoint")
    return model
