"""CPU restatement (functional, torch CPU tensors) of the transformer math on the hot path.

TEST INFRASTRUCTURE (see oracle/__init__.py) — never imported by the product.

PINNED: ``tests/test_oracle_model.py`` checks every function here against golden vectors produced by the
UNMODIFIED reference (``tests/golden/make_golden.py``) and, when /root/reference is present, against the
reference modules directly.  Each function cites the reference lines it restates.

The weights are a plain ``dict[str, Tensor]`` with the reference's state-dict keys
(``transformer.h.{i}.attn.attn.weight`` ...), so a golden fixture is just a state dict + inputs + logits.
Computation dtype = dtype of the tensors passed in; rounding points follow the reference op by op, so running
this with bf16 tensors reproduces the reference's bf16-true CPU numerics, and fp32 gives the wide reference.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor


@dataclass
class RefConfig:
    """The subset of lit_gpt/config.py:16-89 the forward reads."""

    n_layer: int
    n_head: int
    n_embd: int
    n_query_groups: int
    head_size: int
    rope_n_elem: int
    intermediate_size: int
    padded_vocab_size: int
    block_size: int = 4096
    norm_eps: float = 1e-5
    rope_base: int = 10000
    rope_condense_ratio: int = 1
    parallel_residual: bool = False
    shared_attention_norm: bool = False
    norm_class: str = "RMSNorm"  # or "LayerNorm"
    mlp_class: str = "LLaMAMLP"  # or "GptNeoxMLP"
    gelu_approximate: str = "none"

    @classmethod
    def from_any(cls, cfg) -> "RefConfig":
        """From a lit_gpt Config (reference's or ours) or a dict with the same attribute names."""
        get = (lambda k: cfg[k]) if isinstance(cfg, dict) else (lambda k: getattr(cfg, k))
        return cls(
            n_layer=get("n_layer"), n_head=get("n_head"), n_embd=get("n_embd"),
            n_query_groups=get("n_query_groups"), head_size=get("head_size"), rope_n_elem=get("rope_n_elem"),
            intermediate_size=get("intermediate_size"), padded_vocab_size=get("padded_vocab_size"),
            block_size=get("block_size"), norm_eps=get("norm_eps"), rope_base=get("rope_base"),
            rope_condense_ratio=get("rope_condense_ratio"), parallel_residual=get("parallel_residual"),
            shared_attention_norm=get("shared_attention_norm"), norm_class=get("_norm_class"),
            mlp_class=get("_mlp_class"), gelu_approximate=get("gelu_approximate"),
        )


# -- lit_gpt/rmsnorm.py:19-25 -----------------------------------------------------------------------
def rmsnorm(x: Tensor, weight: Tensor, eps: float) -> Tensor:
    dtype = x.dtype
    xf = x.float()
    norm_x = torch.mean(xf * xf, dim=-1, keepdim=True)
    x_normed = xf * torch.rsqrt(norm_x + eps)
    return (weight * x_normed).to(dtype=dtype)  # weight promoted to fp32; single cast at the end


def layernorm(x: Tensor, weight: Tensor, bias: Optional[Tensor], eps: float) -> Tensor:
    return F.layer_norm(x, (x.size(-1),), weight, bias, eps)


# -- lit_gpt/model.py:746-764 -----------------------------------------------------------------------
def build_rope_cache(seq_len: int, n_elem: int, base: int = 10000, condense_ratio: int = 1,
                     pos_dtype: Optional[torch.dtype] = None) -> Tuple[Tensor, Tensor]:
    """`pos_dtype` restates SURVEY F5: the reference divides an int64 arange by a python int, which yields
    the DEFAULT float dtype (bf16 under `--precision bf16-true`), so positions alias above 256."""
    theta = 1.0 / (base ** (torch.arange(0, n_elem, 2).float() / n_elem))
    seq_idx = torch.arange(seq_len).to(pos_dtype or torch.get_default_dtype()) / condense_ratio
    idx_theta = torch.outer(seq_idx, theta).repeat(1, 2)
    return torch.cos(idx_theta), torch.sin(idx_theta)


# -- lit_gpt/model.py:767-773 -----------------------------------------------------------------------
def apply_rope(x: Tensor, cos: Tensor, sin: Tensor) -> Tensor:
    hs = x.size(-1)
    x1, x2 = x[..., : hs // 2], x[..., hs // 2 :]
    rotated = torch.cat((-x2, x1), dim=-1)
    roped = (x * cos) + (rotated * sin)  # promotes to the wider of x / table dtype
    return roped.to(dtype=x.dtype)


def linear(x: Tensor, w: Tensor, b: Optional[Tensor] = None) -> Tensor:
    return F.linear(x, w, b)


class RefKVCache:
    """lit_gpt/model.py:776-799 + build_kv_cache 667-688: buffers of `heads` = n_head (1 when MQA)."""

    def __init__(self, cfg: RefConfig, batch: int, max_seq: int, rope_cache_length: int, dtype: torch.dtype):
        heads = 1 if cfg.n_query_groups == 1 else cfg.n_head
        self.k = torch.zeros(batch, heads, max_seq, rope_cache_length + cfg.head_size - cfg.rope_n_elem, dtype=dtype)
        self.v = torch.zeros(batch, heads, max_seq, cfg.head_size, dtype=dtype)

    def update(self, input_pos: Tensor, k: Tensor, v: Tensor) -> Tuple[Tensor, Tensor]:
        self.k = self.k.to(k.dtype)
        self.v = self.v.to(v.dtype)
        return self.k.index_copy_(2, input_pos, k), self.v.index_copy_(2, input_pos, v)


# -- lit_gpt/model.py:609-665 -----------------------------------------------------------------------
def attention(cfg: RefConfig, p: Dict[str, Tensor], prefix: str, x: Tensor, cos: Tensor, sin: Tensor,
              mask: Optional[Tensor], input_pos: Optional[Tensor], kv: Optional[RefKVCache]) -> Tensor:
    B, T, _ = x.size()
    qkv = linear(x, p[prefix + "attn.weight"], p.get(prefix + "attn.bias"))
    q_per_kv = cfg.n_head // cfg.n_query_groups
    total_qkv = q_per_kv + 2
    qkv = qkv.view(B, T, cfg.n_query_groups, total_qkv, cfg.head_size).permute(0, 2, 3, 1, 4)
    q, k, v = qkv.split((q_per_kv, 1, 1), dim=2)
    if cfg.n_query_groups != cfg.n_head and (input_pos is None or cfg.n_query_groups != 1):
        k = k.expand(B, cfg.n_query_groups, q_per_kv, T, cfg.head_size)
        v = v.expand(B, cfg.n_query_groups, q_per_kv, T, cfg.head_size)
    q = q.reshape(B, -1, T, cfg.head_size)
    k = k.reshape(B, -1, T, cfg.head_size)
    v = v.reshape(B, -1, T, cfg.head_size)
    n = cfg.rope_n_elem
    q = torch.cat((apply_rope(q[..., :n], cos, sin), q[..., n:]), dim=-1)
    k = torch.cat((apply_rope(k[..., :n], cos, sin), k[..., n:]), dim=-1)
    if input_pos is not None:
        if kv is None:
            raise TypeError("You need to call `gpt.set_kv_cache()`")
        k, v = kv.update(input_pos, k, v)
    scale = 1.0 / math.sqrt(cfg.head_size)
    y = F.scaled_dot_product_attention(q, k, v, attn_mask=mask, dropout_p=0.0, scale=scale, is_causal=mask is None)
    y = y.transpose(1, 2).reshape(B, T, cfg.n_head * cfg.head_size)
    return linear(y, p[prefix + "proj.weight"], p.get(prefix + "proj.bias"))


# -- lit_gpt/model.py:691-716 -----------------------------------------------------------------------
def mlp(cfg: RefConfig, p: Dict[str, Tensor], prefix: str, x: Tensor) -> Tensor:
    if cfg.mlp_class == "LLaMAMLP":
        a = linear(x, p[prefix + "fc_1.weight"], p.get(prefix + "fc_1.bias"))
        b = linear(x, p[prefix + "fc_2.weight"], p.get(prefix + "fc_2.bias"))
        return linear(F.silu(a) * b, p[prefix + "proj.weight"], p.get(prefix + "proj.bias"))
    h = linear(x, p[prefix + "fc.weight"], p.get(prefix + "fc.bias"))
    h = F.gelu(h, approximate=cfg.gelu_approximate)
    return linear(h, p[prefix + "proj.weight"], p.get(prefix + "proj.bias"))


def _norm(cfg: RefConfig, p: Dict[str, Tensor], prefix: str, x: Tensor) -> Tensor:
    if cfg.norm_class == "RMSNorm":
        return rmsnorm(x, p[prefix + "weight"], cfg.norm_eps)
    return layernorm(x, p[prefix + "weight"], p.get(prefix + "bias"), cfg.norm_eps)


# -- lit_gpt/model.py:572-593 -----------------------------------------------------------------------
def block(cfg: RefConfig, p: Dict[str, Tensor], i: int, x: Tensor, cos, sin, mask, input_pos, kv) -> Tensor:
    pre = f"transformer.h.{i}."
    n_1 = _norm(cfg, p, pre + "norm_1.", x)
    h = attention(cfg, p, pre + "attn.", n_1, cos, sin, mask, input_pos, kv)
    if cfg.parallel_residual:
        n_2 = n_1 if cfg.shared_attention_norm else _norm(cfg, p, pre + "norm_2.", x)
        return mlp(cfg, p, pre + "mlp.", n_2) + h + x
    x = h + x
    return mlp(cfg, p, pre + "mlp.", _norm(cfg, p, pre + "norm_2.", x)) + x


class RefState:
    """max_seq_length setter + set_kv_cache (lit_gpt/model.py:466-484, 534-554, 802-804)."""

    def __init__(self, cfg: RefConfig, max_seq_length: int, batch: int, dtype: torch.dtype,
                 cos: Optional[Tensor] = None, sin: Optional[Tensor] = None, pos_dtype: Optional[torch.dtype] = None):
        if max_seq_length > cfg.block_size:
            raise ValueError(f"Cannot attend to {max_seq_length}, block size is only {cfg.block_size}")
        self.max_seq_length = max_seq_length
        if cos is None:
            cos, sin = build_rope_cache(max_seq_length, cfg.rope_n_elem, cfg.rope_base, cfg.rope_condense_ratio, pos_dtype)
        self.cos, self.sin = cos, sin
        self.kv: List[RefKVCache] = [RefKVCache(cfg, batch, max_seq_length, cos.size(-1), dtype) for _ in range(cfg.n_layer)]
        self.mask_cache = torch.tril(torch.ones(max_seq_length, max_seq_length, dtype=torch.bool))[None, None]


# -- lit_gpt/model.py:499-519 -----------------------------------------------------------------------
def gpt_forward(cfg: RefConfig, p: Dict[str, Tensor], idx: Tensor, input_pos: Optional[Tensor] = None,
                state: Optional[RefState] = None, cos: Optional[Tensor] = None, sin: Optional[Tensor] = None) -> Tensor:
    T = idx.size(1)
    if state is not None:
        if state.max_seq_length < T:
            raise ValueError(f"Cannot forward sequence of length {T}, max seq length is only {state.max_seq_length}.")
        cos, sin = state.cos, state.sin
    if input_pos is not None:
        cos_t, sin_t = cos.index_select(0, input_pos), sin.index_select(0, input_pos)
        mask = state.mask_cache.index_select(2, input_pos)
    else:
        cos_t, sin_t, mask = cos[:T], sin[:T], None
    x = F.embedding(idx, p["transformer.wte.weight"])
    for i in range(cfg.n_layer):
        kv = state.kv[i] if (state is not None and input_pos is not None) else None
        x = block(cfg, p, i, x, cos_t, sin_t, mask, input_pos, kv)
    x = _norm(cfg, p, "transformer.ln_f.", x)
    return linear(x, p["lm_head.weight"], p.get("lm_head.bias"))


# -- generate/base.py:30-93 -------------------------------------------------------------------------
def sample(logits: Tensor, temperature: float = 1.0, top_k: Optional[int] = None) -> Tensor:
    logits = logits[0, -1]
    if top_k is not None:
        v, i = torch.topk(logits, min(top_k, logits.size(-1)))
        logits = torch.full_like(logits, float("-inf")).scatter_(-1, i, v)
    if temperature > 0.0:
        probs = F.softmax(logits / temperature, dim=-1)
        return torch.multinomial(probs, num_samples=1)
    return torch.argmax(logits, dim=-1, keepdim=True)


@torch.inference_mode()
def generate(cfg: RefConfig, p: Dict[str, Tensor], state: RefState, prompt: Tensor, max_returned_tokens: int, *,
             temperature: float = 1.0, top_k: Optional[int] = None, eos_id: Optional[int] = None,
             return_logits: bool = False, forced: Optional[Tensor] = None):
    """`forced` (test helper, no counterpart in the reference): a full token sequence to follow instead of sampling, so that
    two precisions can be compared step by step along the SAME tokens (teacher forcing)."""
    T = prompt.size(0)
    assert max_returned_tokens > T
    if state.max_seq_length < max_returned_tokens - 1:
        raise NotImplementedError(f"max_seq_length {state.max_seq_length} needs to be >= {max_returned_tokens - 1}")
    tokens = [prompt]
    all_logits = []
    input_pos = torch.tensor([T])
    logits = gpt_forward(cfg, p, prompt.view(1, -1), torch.arange(0, T), state)
    all_logits.append(logits[0, -1].clone())
    token = sample(logits, temperature, top_k).to(prompt.dtype) if forced is None else forced[T].view(1).to(prompt.dtype)
    tokens.append(token)
    for i in range(2, max_returned_tokens - T + 1):
        logits = gpt_forward(cfg, p, token.view(1, -1), input_pos, state)
        all_logits.append(logits[0, -1].clone())
        token = sample(logits, temperature, top_k).to(prompt.dtype) if forced is None else forced[T + i - 1].view(1).to(prompt.dtype)
        tokens.append(token)
        if eos_id is not None and token == eos_id:
            break
        input_pos = input_pos.add_(1)
    out = torch.cat(tokens)
    return (out, torch.stack(all_logits)) if return_logits else out


# -- generate/tp.py:28-45 ---------------------------------------------------------------------------
def tensor_parallel_shard(weight: Tensor, world_size: int, rank: int, style: str) -> Tensor:
    dim = {"colwise": 0, "rowwise": 1}[style]
    if weight.size(dim) % world_size != 0:
        raise ValueError("not evenly divisible by the world size")
    return torch.tensor_split(weight, world_size, dim=dim)[rank]


def init_weights(cfg: RefConfig, seed: int, dtype: torch.dtype = torch.float32, std: float = 0.02,
                 bias: bool = False) -> Dict[str, Tensor]:
    """Random-init weights with the reference's parameter names/shapes (GPT.__init__, model.py:446-460;
    init law model.py:490-497: N(0, 0.02) Linears/Embedding, ones for norm weights)."""
    g = torch.Generator().manual_seed(seed)
    p: Dict[str, Tensor] = {}

    def lin(name, out_f, in_f, with_bias=bias):
        p[name + ".weight"] = (torch.randn(out_f, in_f, generator=g) * std).to(dtype)
        if with_bias:
            p[name + ".bias"] = (torch.randn(out_f, generator=g) * std).to(dtype)

    def norm(name):
        p[name + ".weight"] = torch.ones(cfg.n_embd, dtype=dtype)
        if cfg.norm_class == "LayerNorm":
            p[name + ".bias"] = torch.zeros(cfg.n_embd, dtype=dtype)

    lin("lm_head", cfg.padded_vocab_size, cfg.n_embd, False)
    p["transformer.wte.weight"] = (torch.randn(cfg.padded_vocab_size, cfg.n_embd, generator=g) * std).to(dtype)
    for i in range(cfg.n_layer):
        pre = f"transformer.h.{i}."
        norm(pre + "norm_1")
        lin(pre + "attn.attn", (cfg.n_head + 2 * cfg.n_query_groups) * cfg.head_size, cfg.n_embd)
        lin(pre + "attn.proj", cfg.n_embd, cfg.n_embd)
        if not cfg.shared_attention_norm:
            norm(pre + "norm_2")
        if cfg.mlp_class == "LLaMAMLP":
            lin(pre + "mlp.fc_1", cfg.intermediate_size, cfg.n_embd)
            lin(pre + "mlp.fc_2", cfg.intermediate_size, cfg.n_embd)
        else:
            lin(pre + "mlp.fc", cfg.intermediate_size, cfg.n_embd)
        lin(pre + "mlp.proj", cfg.n_embd, cfg.intermediate_size)
    norm("transformer.ln_f")
    return p
