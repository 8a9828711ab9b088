"""``torch.library`` custom ops over the C-ABI kernels (SURVEY §8b "Torch custom ops").

The reference has no custom ops: its quantized Linear, RMSNorm and attention are compositions of ATen ops that
``torch.compile`` traces (tests/test_model.py:564-575 asserts zero graph breaks).  The replacements are opaque kernels,
so each one is registered here with a schema and a fake (meta) kernel; a traced ``GPT.forward`` then sees single
``b200gpt::*`` nodes instead of graph breaks.  Ops that write the KV cache declare it in ``mutates_args``.

All ops take and return plain tensors; the real implementations are the ctypes wrappers in ``kernels.py`` (which launch on
the current stream and never synchronise), so the ops are CUDA-graph capturable.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import kernels as K

NS = "b200gpt"


# ---- quantized Linear: y = x @ dequant(W).T (+ bias); bnb.nn.Linear4bit.forward at model.py:519,619,656,713,714,716 ----
@torch.library.custom_op(f"{NS}::qlinear", mutates_args=())
def qlinear(x: torch.Tensor, packed: torch.Tensor, scale: torch.Tensor, out_features: int, in_features: int, quant_type: str,
            blocksize: int, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    rows = x.numel() // in_features
    if rows <= 16:
        return K.qgemv(x, packed, scale, out_features, in_features, quant_type, blocksize, bias=bias)
    return K.qgemm(x, packed, scale, out_features, in_features, quant_type, blocksize, bias=bias)


@qlinear.register_fake
def _(x, packed, scale, out_features, in_features, quant_type, blocksize, bias=None):
    return x.new_empty((*x.shape[:-1], out_features), dtype=torch.bfloat16)


@torch.library.custom_op(f"{NS}::qlinear_tiled", mutates_args=())
def qlinear_tiled(x: torch.Tensor, tiled: torch.Tensor, out_features: int, in_features: int, quant_type: str, blocksize: int,
                  bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """<= 16 rows over the tiled layout (csrc/tile.cu): the streaming decode GEMV."""
    return K.qgemv_tiled(x, tiled, out_features, in_features, quant_type, blocksize, bias=bias)


@qlinear_tiled.register_fake
def _(x, tiled, out_features, in_features, quant_type, blocksize, bias=None):
    return x.new_empty((*x.shape[:-1], out_features), dtype=torch.bfloat16)


# ---- RMSNorm (lit_gpt/rmsnorm.py:19-25) -------------------------------------------------------------------------
@torch.library.custom_op(f"{NS}::rmsnorm", mutates_args=())
def rmsnorm(x: torch.Tensor, weight: torch.Tensor, eps: float) -> torch.Tensor:
    return K.rmsnorm(x, weight, eps)


@rmsnorm.register_fake
def _(x, weight, eps):
    return torch.empty_like(x)


# ---- decode attention: RoPE + cache append + SDPA for one token per sequence (model.py:624-653) -------------------
@torch.library.custom_op(f"{NS}::attn_decode", mutates_args=("k_cache", "v_cache"))
def attn_decode(qkv: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, pos: torch.Tensor, k_cache: torch.Tensor,
                v_cache: torch.Tensor, n_head: int, n_groups: int, head_size: int, rope_n_elem: int,
                table_preselected: bool = False) -> torch.Tensor:
    return K.attn_decode(qkv, cos, sin, pos, k_cache, v_cache, n_head, n_groups, head_size, rope_n_elem,
                         table_preselected=table_preselected)


@attn_decode.register_fake
def _(qkv, cos, sin, pos, k_cache, v_cache, n_head, n_groups, head_size, rope_n_elem, table_preselected=False):
    return qkv.new_empty((k_cache.shape[0], n_head * head_size), dtype=torch.bfloat16)


# ---- prefill: RoPE + cache write, then causal attention over the cache (model.py:624-665) -------------------------
@torch.library.custom_op(f"{NS}::rope_kv_append", mutates_args=("k_cache", "v_cache"))
def rope_kv_append(qkv: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, pos: torch.Tensor, k_cache: torch.Tensor,
                   v_cache: torch.Tensor, n_head: int, n_groups: int, head_size: int, rope_n_elem: int,
                   table_preselected: bool = False) -> torch.Tensor:
    return K.rope_kv_append(qkv, cos, sin, pos, k_cache, v_cache, n_head, n_groups, head_size, rope_n_elem,
                            table_preselected=table_preselected)


@rope_kv_append.register_fake
def _(qkv, cos, sin, pos, k_cache, v_cache, n_head, n_groups, head_size, rope_n_elem, table_preselected=False):
    return qkv.new_empty((qkv.shape[0], n_head, qkv.shape[1], head_size), dtype=torch.bfloat16)


@torch.library.custom_op(f"{NS}::attn_prefill", mutates_args=())
def attn_prefill(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, pos: torch.Tensor, scale: float) -> torch.Tensor:
    return K.attn_prefill(q, k_cache, v_cache, pos, scale)


@attn_prefill.register_fake
def _(q, k_cache, v_cache, pos, scale):
    B, H, T, hs = q.shape
    return q.new_empty((B, T, H * hs), dtype=torch.bfloat16)


# ---- sampling: logits[0, -1] -> one token (generate/base.py:22-41) -----------------------------------------------
@torch.library.custom_op(f"{NS}::sample", mutates_args=("rng_state",))
def sample(logits: torch.Tensor, temperature: float, top_k: int, rng_state: torch.Tensor) -> torch.Tensor:
    """top_k <= 0: no truncation.  rng_state: int64[2] = (seed, offset) on the device, advanced by the kernel."""
    out = torch.empty(1, dtype=torch.int64, device=logits.device)
    K.sample_(logits, out, temperature, top_k if top_k > 0 else None, rng_state)
    return out


@sample.register_fake
def _(logits, temperature, top_k, rng_state):
    return logits.new_empty((1,), dtype=torch.int64)
