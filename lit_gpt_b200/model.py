"""Decoder-only transformer with the reference's public surface (``lit_gpt/model.py:445-804``): ``GPT``, ``Block``,
``CausalSelfAttention``, ``GptNeoxMLP``, ``LLaMAMLP``, ``KVCache``, ``build_rope_cache``, ``apply_rope``,
``build_mask_cache`` — same constructor signatures, parameter names (state-dict compatible) and error behaviour.

What differs is what runs underneath on a B200:

* CUDA + bf16: attention goes through the library's fused kernels (RoPE + KV append + split-K flash-decode for T = 1,
  RoPE + KV append then SDPA for prefill) over a GQA-NATIVE cache (``n_query_groups`` heads, not ``n_head``).
* When every Linear is a quantized ``QLinear`` and the block is Llama-shaped, a decode step (T = 1) does not walk the
  module tree at all: ``GPT._fused_decode`` issues 5 kernels per block (norm folded into GEMV prologues, residual and
  SwiGLU into epilogues), all capturable in one CUDA graph.
* Anything else (fp32 / CPU, e.g. BASELINE config 1) runs the reference formulation in plain torch, so results are
  identical to the reference there.
"""
from __future__ import annotations

import math
import os
from typing import Any, List, Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import kernels as K
from .config import Config
from .quant import QLinear

# batched decode with more rows than this runs RMSNorm as its own launch instead of in the GEMV prologue (see GPT._fused_decode)
SPLIT_NORM_ROWS = int(os.environ.get("B200_SPLIT_NORM_ROWS", "8"))

FUSED_HEAD_SIZES = (64, 128)


class GPT(nn.Module):
    def __init__(self, config: Config) -> None:
        super().__init__()
        assert config.padded_vocab_size is not None
        self.config = config
        self.lm_head = nn.Linear(config.n_embd, config.padded_vocab_size, bias=config.lm_head_bias)
        self.transformer = nn.ModuleDict(
            dict(
                wte=nn.Embedding(config.padded_vocab_size, config.n_embd),
                h=nn.ModuleList(Block(config) for _ in range(config.n_layer)),
                ln_f=config.norm_class(config.n_embd, eps=config.norm_eps),
            )
        )
        self.max_seq_length = self.config.block_size
        self.mask_cache: Optional[torch.Tensor] = None
        # tensor-parallel state, set by generate.tp.tensor_parallel (None = single GPU)
        self.tp_group: Any = None
        self.tp_world_size: int = 1

    @property
    def max_seq_length(self) -> int:
        return self._max_seq_length

    @max_seq_length.setter
    def max_seq_length(self, value: int) -> None:
        """Smaller than block_size to bound cache memory (reference model.py:466-484)."""
        if value > self.config.block_size:
            raise ValueError(f"Cannot attend to {value}, block size is only {self.config.block_size}")
        self._max_seq_length = value
        if not hasattr(self, "cos"):
            cos, sin = self.rope_cache()
            self.register_buffer("cos", cos, persistent=False)
            self.register_buffer("sin", sin, persistent=False)
        elif value != self.cos.size(0):
            self.cos, self.sin = self.rope_cache(device=self.cos.device)

    def reset_parameters(self) -> None:
        self.cos, self.sin = self.rope_cache()

    def _init_weights(self, module: nn.Module) -> None:
        """`gpt.apply(gpt._init_weights)`: N(0, 0.02) for Linear / Embedding weights, zero biases."""
        if isinstance(module, (nn.Linear, QLinear)):
            if module.weight.is_floating_point():
                torch.nn.init.normal_(module.weight, mean=0.0, std=0.02)
            if module.bias is not None:
                torch.nn.init.zeros_(module.bias)
        elif isinstance(module, nn.Embedding):
            torch.nn.init.normal_(module.weight, mean=0.0, std=0.02)

    def forward(self, idx: torch.Tensor, input_pos: Optional[torch.Tensor] = None) -> torch.Tensor:
        T = idx.size(1)
        if self.max_seq_length < T:
            raise ValueError(f"Cannot forward sequence of length {T}, max seq length is only {self.max_seq_length}.")

        if input_pos is not None:
            if self.mask_cache is None:
                raise TypeError("You need to call `gpt.set_kv_cache()`")
            # (a traced forward takes the module path: its kernels are registered custom ops, the fused chain is raw C-ABI calls)
            if T == 1 and input_pos.numel() == 1 and not torch.compiler.is_compiling() and self.fused_decode_ready(idx):
                return self._fused_decode(idx, input_pos)
            if (T > 1 and input_pos.numel() == T and input_pos.is_cuda and not torch.compiler.is_compiling() and self.fused_decode_ready(idx)
                    and self.config.intermediate_size % 8 == 0):
                return self._fused_prefill(idx, input_pos)
            cos = self.cos.index_select(0, input_pos)
            sin = self.sin.index_select(0, input_pos)
            mask = self.mask_cache.index_select(2, input_pos)
        else:
            cos = self.cos[:T]
            sin = self.sin[:T]
            mask = None

        x = self.transformer.wte(idx)
        for block in self.transformer.h:
            x = block(x, cos, sin, mask, input_pos)
        x = self.transformer.ln_f(x)
        return self.lm_head(x)

    @classmethod
    def from_name(cls, name: str, **kwargs: Any) -> "GPT":
        return cls(Config.from_name(name, **kwargs))

    def rope_cache(self, device: Optional[torch.device] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        return build_rope_cache(
            seq_len=self.max_seq_length,
            n_elem=self.config.rope_n_elem,
            device=device,
            condense_ratio=self.config.rope_condense_ratio,
            base=self.config.rope_base,
        )

    def set_kv_cache(self, batch_size: int, rope_cache_length: Optional[int] = None,
                     device: Optional[torch.device] = None, dtype: Optional[torch.dtype] = None) -> None:
        if rope_cache_length is None:
            rope_cache_length = self.cos.size(-1)
        max_seq_length = self.max_seq_length
        for block in self.transformer.h:
            block.attn.kv_cache = block.attn.build_kv_cache(batch_size, max_seq_length, rope_cache_length, device, dtype)
        if self.mask_cache is None or self.mask_cache.size(3) != max_seq_length:
            self.mask_cache = build_mask_cache(max_seq_length, device)
        self._fused_plan = None
        self.__dict__.pop("_decode_graphs", None)  # captured graphs pin the kv / table / weight buffers
        self.__dict__.pop("_prefill_graphs", None)

    def clear_kv_cache(self) -> None:
        self.mask_cache = None
        for block in self.transformer.h:
            block.attn.kv_cache = None
        self._fused_plan = None
        self.__dict__.pop("_decode_graphs", None)  # captured graphs pin the kv / table / weight buffers
        self.__dict__.pop("_prefill_graphs", None)

    # ------------------------------------------------------------------------------------------------------
    # fused decode step
    # ------------------------------------------------------------------------------------------------------
    _fused_plan: Optional[bool] = None

    def _apply(self, fn, recurse: bool = True):
        self._fused_plan = None  # device moves quantize the Linears: re-plan on the next forward
        self.__dict__.pop("_spans", None)
        self.__dict__.pop("_seq_tables", None)
        self.__dict__.pop("_p2p", None)
        self.__dict__.pop("_decode_graphs", None)  # captured graphs pin the kv / table / weight buffers
        self.__dict__.pop("_prefill_graphs", None)
        return super()._apply(fn, recurse)

    def fused_decode_ready(self, idx: Optional[torch.Tensor] = None) -> bool:
        """True when a T=1 step can run as the 5-kernels-per-block fused chain (see module docstring)."""
        if idx is not None and not idx.is_cuda:
            return False
        if self._fused_plan is None:
            self._fused_plan = self._check_fused()
        return self._fused_plan

    def _check_fused(self) -> bool:
        c = self.config
        if c._norm_class != "RMSNorm" or c._mlp_class != "LLaMAMLP" or c.parallel_residual or c.shared_attention_norm:
            return False
        if c.n_expert or c.head_size not in FUSED_HEAD_SIZES or c.rope_n_elem % 2:
            return False
        wte = self.transformer.wte.weight
        if not wte.is_cuda or wte.dtype != torch.bfloat16:
            return False

        def ok(lin: nn.Module) -> bool:
            return (isinstance(lin, QLinear) and lin.is_quantized and lin.weight.is_cuda and lin.bias is None
                    and lin.quant_type != "int8")

        for block in self.transformer.h:
            a, m = block.attn, block.mlp
            if not (ok(a.attn) and ok(a.proj) and ok(m.fc_1) and ok(m.fc_2) and ok(m.proj)):
                return False
            if m.fc_1.quant_type != m.fc_2.quant_type:
                return False
            if not isinstance(a.kv_cache, KVCache) or not a.kv_cache.gqa_native:
                return False
            if block.norm_1.weight.dtype != torch.bfloat16 or block.norm_2.weight.dtype != torch.bfloat16:
                return False
            # forward hooks other than our own TP marker would be bypassed by the fused chain
            if (a._forward_hooks or m._forward_hooks) and self.tp_world_size == 1:
                return False
        return ok(self.lm_head) and self.transformer.ln_f.weight.dtype == torch.bfloat16

    def _all_reduce(self, t: torch.Tensor) -> torch.Tensor:
        import torch.distributed as dist

        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.tp_group)  # NCCL over NVLink; sums row-parallel partials
        return t

    # ---- tensor-parallel hand-off over peer memory (lit_gpt_b200/p2p.py) instead of NCCL on the decode path ----------
    def _tp_comm(self, batch: int) -> Optional[Any]:
        """The peer-memory communicator of the fused decode step, created collectively on first use (never during graph
        capture); None -> the NCCL all-reduce path (B200_TP_NCCL=1, non-tiled shapes, or peer mapping unavailable)."""
        st = self.__dict__.get("_p2p")
        if st is not None:
            return st or None
        if torch.cuda.is_current_stream_capturing():
            return None

        from .p2p import PeerComm

        blocks = list(self.transformer.h)
        # mlp.proj may lack a tiled layout (row-parallel shard whose in_features is not a multiple of the quantisation block,
        # e.g. 11008 / 8): that hand-off alone then stays an NCCL all-reduce, the attention branch still uses peer memory
        lins = [self.lm_head] + [l for b in blocks for l in (b.attn.attn, b.attn.proj, b.mlp.fc_1, b.mlp.fc_2)]
        lins += [b.mlp.proj for b in blocks if b.mlp.proj.tiled is not None]
        ok = os.environ.get("B200_TP_NCCL", "0") != "1" and all(l.streams(batch) for l in lins)
        comm: Any = False
        if ok:
            try:
                d_full = blocks[0].attn.proj.out_features
                gather = 16 * self.lm_head.out_features * 4 if getattr(self, "lm_head_sharded", False) else 0
                comm = PeerComm(self.tp_group, self.transformer.wte.weight.device, 16 * d_full * 4, gather)  # 32-bit packets
            except Exception as e:  # consistent on all ranks (PeerComm agrees on failure collectively)
                import warnings

                warnings.warn(f"peer-memory hand-off unavailable ({e}); tensor-parallel decode uses NCCL all-reduce")
                comm = False
        self.__dict__["_p2p"] = comm
        return comm or None

    # ---- blocks placed on several devices by generate/sequentially.py ----------------------------------------------------
    def release_reference_layouts(self) -> int:
        """Keep ONE resident copy of every 4-bit Linear (the tiled layout the decode GEMV streams); see
        ``QLinear.release_reference_layout``.  Returns the number of Linears whose reference-layout copy was dropped."""
        n = sum(int(m.release_reference_layout()) for m in self.modules() if isinstance(m, QLinear))
        self._fused_plan = None
        return n

    def spans_devices(self) -> bool:
        hit = self.__dict__.get("_spans")
        if hit is None:
            hit = self.__dict__["_spans"] = len({b.norm_1.weight.device for b in self.transformer.h}) > 1
        return hit

    def _tables_on(self, dev: torch.device) -> Tuple[torch.Tensor, torch.Tensor]:
        """cos / sin on `dev` (every device of a sequentially placed model keeps its own copy of the RoPE tables)."""
        if dev == self.cos.device:
            return self.cos, self.sin
        tabs = self.__dict__.setdefault("_seq_tables", {})
        hit = tabs.get(dev)
        if hit is None or hit[2] != self.cos.data_ptr():
            hit = tabs[dev] = (self.cos.to(dev), self.sin.to(dev), self.cos.data_ptr())
        return hit[0], hit[1]

    def _fused_prefill(self, idx: torch.Tensor, input_pos: torch.Tensor) -> torch.Tensor:
        """T > 1 tokens over the cache with this library's kernels only (reference: model.py:499-519, 572-593, 609-665, 712-716): the
        embedding, the residual adds fused into the RMSNorm that follows them (b200_rmsnorm delta / sum_out), SwiGLU as one kernel,
        RoPE + cache append + causal attention straight from the full cos / sin tables and input_pos -- no table or mask
        `index_select` (the 16 MB mask gather at T = 4096 fed nothing: the attention kernel derives the mask row from input_pos)."""
        c = self.config
        B, T = idx.shape
        eps = c.norm_eps
        tp = self.tp_world_size > 1
        scale = 1.0 / math.sqrt(c.head_size)
        if self.spans_devices():
            return self._fused_prefill_multi(idx, input_pos)
        x = K.embedding(idx.reshape(B * T), self.transformer.wte.weight)            # (B*T, d)
        delta: Optional[torch.Tensor] = None
        for block in self.transformer.h:
            a, m, cache = block.attn, block.mlp, block.attn.kv_cache
            if delta is None:
                n1 = K.rmsnorm(x, block.norm_1.weight, eps)
            else:                                                                      # x = x + mlp output, then norm_1
                x_new = torch.empty_like(x)
                n1 = K.rmsnorm(x, block.norm_1.weight, eps, delta=delta, sum_out=x_new)
                x = x_new
            qkv = a.attn(n1.view(B, T, -1))
            q = K.rope_kv_append(qkv, self.cos, self.sin, input_pos, cache.k, cache.v, c.n_head, c.n_query_groups, c.head_size,
                                 c.rope_n_elem, table_preselected=False)
            y = K.attn_prefill(q, cache.k, cache.v, input_pos, scale)
            h = a.proj(y)
            if tp:
                h = self._all_reduce(h.contiguous())
            x_new = torch.empty_like(x)
            n2 = K.rmsnorm(x, block.norm_2.weight, eps, delta=h.view(B * T, -1), sum_out=x_new)   # x = h + x, then norm_2
            x = x_new
            n2 = n2.view(B, T, -1)
            delta = m(n2).view(B * T, -1)   # LLaMAMLP.forward: fused SwiGLU (b200_swiglu, or the GEMV epilogue for <= 16 rows); its tensor-parallel hook sums the ranks' partials
        xf = K.rmsnorm(x, self.transformer.ln_f.weight, eps, delta=delta)
        return self.lm_head(xf.view(B, T, -1))

    def _fused_prefill_multi(self, idx: torch.Tensor, input_pos: torch.Tensor) -> torch.Tensor:
        """_fused_prefill for blocks placed on several devices (generate/sequentially.py): the residual stream and the pending
        branch output hop to the next device when a block lives elsewhere; every launch runs with its device current."""
        c = self.config
        B, T = idx.shape
        eps, scale = c.norm_eps, 1.0 / math.sqrt(c.head_size)
        root = self.transformer.wte.weight.device
        with torch.cuda.device(root):
            x = K.embedding(idx.reshape(B * T), self.transformer.wte.weight)
        delta: Optional[torch.Tensor] = None
        for block in self.transformer.h:
            dev = block.norm_1.weight.device
            a, m, cache = block.attn, block.mlp, block.attn.kv_cache
            with torch.cuda.device(dev):
                if x.device != dev:
                    x = x.to(dev, non_blocking=True)
                    delta = None if delta is None else delta.to(dev, non_blocking=True)
                cos, sin = self._tables_on(dev)
                pos = input_pos if input_pos.device == dev else input_pos.to(dev, non_blocking=True)
                if delta is None:
                    n1 = K.rmsnorm(x, block.norm_1.weight, eps)
                else:
                    x_new = torch.empty_like(x)
                    n1 = K.rmsnorm(x, block.norm_1.weight, eps, delta=delta, sum_out=x_new)
                    x = x_new
                qkv = a.attn(n1.view(B, T, -1))
                q = K.rope_kv_append(qkv, cos, sin, pos, cache.k, cache.v, c.n_head, c.n_query_groups, c.head_size, c.rope_n_elem,
                                     table_preselected=False)
                h = a.proj(K.attn_prefill(q, cache.k, cache.v, pos, scale))
                x_new = torch.empty_like(x)
                n2 = K.rmsnorm(x, block.norm_2.weight, eps, delta=h.view(B * T, -1), sum_out=x_new)
                x = x_new
                delta = m.proj(K.swiglu(m.fc_1(n2.view(B, T, -1)), m.fc_2(n2.view(B, T, -1)))).view(B * T, -1)
        with torch.cuda.device(root):
            if x.device != root:
                x, delta = x.to(root, non_blocking=True), delta.to(root, non_blocking=True)
            xf = K.rmsnorm(x, self.transformer.ln_f.weight, eps, delta=delta)
            return self.lm_head(xf.view(B, T, -1))

    def _fused_decode_multi(self, idx: torch.Tensor, input_pos: torch.Tensor) -> torch.Tensor:
        """_fused_decode for blocks placed on several devices: the same five launches per block, each device segment on its own
        GPU; the 8 KB residual stream crosses NVLink once per device boundary."""
        c = self.config
        B = idx.size(0)
        eps = c.norm_eps
        root = self.transformer.wte.weight.device
        with torch.cuda.device(root):
            x = K.embedding(idx.view(B), self.transformer.wte.weight)
        for block in self.transformer.h:
            dev = block.norm_1.weight.device
            a, m, cache = block.attn, block.mlp, block.attn.kv_cache
            with torch.cuda.device(dev):
                if x.device != dev:
                    x = x.to(dev, non_blocking=True)
                cos, sin = self._tables_on(dev)
                pos = input_pos if input_pos.device == dev else input_pos.to(dev, non_blocking=True)
                if B > SPLIT_NORM_ROWS:   # see _fused_decode
                    qkv = a.attn.gemv(K.rmsnorm(x, block.norm_1.weight, eps), kv_hint=(cache.k, cache.v, pos))
                else:
                    qkv = a.attn.gemv(x, norm_weight=block.norm_1.weight, norm_eps=eps, kv_hint=(cache.k, cache.v, pos))
                y = K.attn_decode(qkv, cos, sin, pos, cache.k, cache.v, c.n_head, c.n_query_groups, c.head_size, c.rope_n_elem)
                x = a.proj.gemv(y, residual=x)
                if B > SPLIT_NORM_ROWS:
                    act = m.fc_1.gemv(K.rmsnorm(x, block.norm_2.weight, eps), other=m.fc_2)
                else:
                    act = m.fc_1.gemv(x, other=m.fc_2, norm_weight=block.norm_2.weight, norm_eps=eps)
                x = m.proj.gemv(act, residual=x)
        with torch.cuda.device(root):
            if x.device != root:
                x = x.to(root, non_blocking=True)
            if B > SPLIT_NORM_ROWS:
                logits = self.lm_head.gemv(K.rmsnorm(x, self.transformer.ln_f.weight, eps))
            else:
                logits = self.lm_head.gemv(x, norm_weight=self.transformer.ln_f.weight, norm_eps=eps)
        return logits.view(B, 1, -1)

    def _gather_dummy(self, B: int, n: int, device: torch.device) -> torch.Tensor:
        """The sharded lm_head's local slice is also stored locally (y): a persistent scratch row, not a per-step allocation."""
        t = self.__dict__.get("_gather_y")
        if t is None or t.shape != (B, n) or t.device != device:
            t = self.__dict__["_gather_y"] = torch.empty((B, n), dtype=torch.bfloat16, device=device)
        return t

    def _fused_decode(self, idx: torch.Tensor, input_pos: torch.Tensor) -> torch.Tensor:
        """One token per sequence.  Same math as Block.forward / CausalSelfAttention.forward / LLaMAMLP.forward of
        the reference (model.py:572-593, 609-656, 712-716) with identical bf16 rounding points."""
        c = self.config
        B = idx.size(0)
        tp = self.tp_world_size > 1
        if not tp and self.spans_devices():
            return self._fused_decode_multi(idx, input_pos)
        x = K.embedding(idx.view(B), self.transformer.wte.weight)  # (B, d_full)
        delta: Optional[torch.Tensor] = None
        eps = c.norm_eps

        def lin(q: QLinear, inp: torch.Tensor, nxt: Optional[QLinear] = None, **kw: Any) -> torch.Tensor:
            # nxt: the Linear of the NEXT GEMV in the step; its tiled weights are prefetched into L2 by this kernel
            return q.gemv(inp, prefetch=None if nxt is None else nxt.tiled, **kw)

        blocks = list(self.transformer.h)
        comm = self._tp_comm(B) if tp and B <= 16 else None
        if comm is not None:
            # TP over peer memory: the row-parallel GEMV's epilogue writes its partial into every rank's slot, the next
            # GEMV's prologue sums the slots (b200_tp_handoff): no all-reduce launch, no NCCL on the decode path.
            # Hand-offs alternate between buffer 0 (attention branch) and buffer 1 (MLP branch).
            def tlin(q: QLinear, inp: torch.Tensor, other: Optional[QLinear] = None, send: Optional[int] = None,
                     recv: Optional[int] = None, **kw: Any) -> torch.Tensor:
                return K.qgemv_tiled(inp, q.tiled, q.out_features, q.in_features, q.quant_type, q.blocksize,
                                     tiled2=None if other is None else other.tiled, tp=comm.handoff(send, recv), fold=q.fold, **kw)

            # pending: how the previous block's MLP output reaches the next prologue: "p2p" (slots of buffer 1), a tensor
            # (NCCL all-reduced delta; the collective also orders the ranks, so reusing buffer 0 next block stays race-free)
            pending: Any = None
            for li, block in enumerate(blocks):
                a, m, cache = block.attn, block.mlp, block.attn.kv_cache
                if pending is None:
                    qkv = lin(a.attn, x, norm_weight=block.norm_1.weight, norm_eps=eps)
                else:
                    x_new = torch.empty_like(x)
                    if isinstance(pending, str):
                        qkv = tlin(a.attn, x, recv=1, x_out=x_new, norm_weight=block.norm_1.weight, norm_eps=eps)
                    else:
                        qkv = lin(a.attn, x, delta=pending, x_out=x_new, norm_weight=block.norm_1.weight, norm_eps=eps)
                    x = x_new
                y = K.attn_decode(qkv, self.cos, self.sin, input_pos, cache.k, cache.v, c.n_head, c.n_query_groups,
                                  c.head_size, c.rope_n_elem)
                tlin(a.proj, y, send=0)
                x_new = torch.empty_like(x)
                act = tlin(m.fc_1, x, other=m.fc_2, recv=0, x_out=x_new, norm_weight=block.norm_2.weight, norm_eps=eps)
                x = x_new
                if m.proj.tiled is not None:
                    tlin(m.proj, act, send=1)
                    pending = "p2p"
                else:
                    pending = self._all_reduce(lin(m.proj, act))
            head = self.lm_head
            if getattr(self, "lm_head_sharded", False):
                # lm_head sharded by rows (§8f-3): this rank's vocabulary slice goes straight from the GEMV's epilogue into every
                # rank's gather buffer; b200_tp_gather unpacks the full logits (a third buffer: the two branch buffers alternate)
                hs = comm.gather_send()
                if isinstance(pending, str):
                    hr = comm.handoff(None, 1)
                    hs.slots, hs.slot_stride, hs.recv_seq, hs.recv_done, hs.reduced = hr.slots, hr.slot_stride, hr.recv_seq, hr.recv_done, hr.reduced
                    K.qgemv_tiled(x, head.tiled, head.out_features, head.in_features, head.quant_type, head.blocksize, tp=hs,
                                  norm_weight=self.transformer.ln_f.weight, norm_eps=eps, out=self._gather_dummy(B, head.out_features, x.device))
                else:
                    K.qgemv_tiled(x, head.tiled, head.out_features, head.in_features, head.quant_type, head.blocksize, tp=hs, delta=pending,
                                  norm_weight=self.transformer.ln_f.weight, norm_eps=eps, out=self._gather_dummy(B, head.out_features, x.device))
                logits = K.tp_gather(comm.gather_recv(), B, head.out_features, x.device)
            elif isinstance(pending, str):
                logits = tlin(head, x, recv=1, norm_weight=self.transformer.ln_f.weight, norm_eps=eps)
            else:
                logits = lin(head, x, delta=pending, norm_weight=self.transformer.ln_f.weight, norm_eps=eps)
            return logits.view(B, 1, -1)

        a0 = blocks[0].attn.attn if blocks else None
        # EXPERIMENT, off by default (B200_FUSE_ATTN_MAX_SEQ=1024 enables it): the attention step fused into the QKV GEMV's
        # kernel (b200_qgemv_tiled_args::attn).  Correct (tests/test_gpu_kernels.py::test_qgemv_tiled_fused_attention) and 32
        # launches per token fewer, but measured SLOWER on B200 (632 vs 676 tok/s): in-kernel the
        # attention tail runs on one SM per head at the GEMV's register budget, 5.5 - 7.5 us against ~6 us for the stand-alone
        # cluster kernel INCLUDING its launch boundary
        fuse_attn = (not tp and B == 1 and a0 is not None and self.max_seq_length <= int(os.environ.get("B200_FUSE_ATTN_MAX_SEQ", "0"))
                     and a0.fold == 1 and a0.streams(1) and input_pos.numel() == 1
                     and K.qgemv_attn_fusable(a0.out_features, a0.in_features, a0.quant_type, c.n_head, c.n_query_groups, c.head_size,
                                              c.rope_n_elem, self.max_seq_length))
        # more than 8 rows: RMSNorm as its own launch.  Fused, EVERY CTA of the GEMV normalises all the rows again (148 x redundant:
        # at 16 rows two thirds of the launch's instructions); plain rows reach shared memory by one bulk copy each
        split_norm = B > SPLIT_NORM_ROWS and not tp
        for li, block in enumerate(blocks):
            a, m, cache = block.attn, block.mlp, block.attn.kv_cache
            nxt_qkv = blocks[li + 1].attn.attn if li + 1 < len(blocks) else self.lm_head
            hint = (cache.k, cache.v, input_pos)  # L2 hint: the filled K / V rows this block's attention reads next
            if fuse_attn:
                # the attention step runs inside the QKV GEMV's kernel (b200_qgemv_tiled_args::attn): one launch boundary less
                y = torch.empty((B, c.n_head * c.head_size), dtype=torch.bfloat16, device=x.device)
                qkv = lin(a.attn, x, a.proj, norm_weight=block.norm_1.weight, norm_eps=eps, kv_hint=hint,
                          attn=dict(k_cache=cache.k, v_cache=cache.v, pos=input_pos, cos=self.cos, sin=self.sin, out=y, n_head=c.n_head,
                                    n_groups=c.n_query_groups, head_size=c.head_size, rope_n_elem=c.rope_n_elem))
            elif delta is not None:  # TP: the all-reduced branch output is added in the next prologue
                x_new = torch.empty_like(x)
                qkv = lin(a.attn, x, a.proj, delta=delta, x_out=x_new, norm_weight=block.norm_1.weight, norm_eps=eps, kv_hint=hint)
                x = x_new
            elif split_norm:
                qkv = lin(a.attn, K.rmsnorm(x, block.norm_1.weight, eps), a.proj, kv_hint=hint)
            else:
                qkv = lin(a.attn, x, a.proj, norm_weight=block.norm_1.weight, norm_eps=eps, kv_hint=hint)
            if not fuse_attn:
                y = K.attn_decode(qkv, self.cos, self.sin, input_pos, cache.k, cache.v, c.n_head, c.n_query_groups,
                                  c.head_size, c.rope_n_elem)
            if tp:
                h = self._all_reduce(lin(a.proj, y, m.fc_1))
                x_new = torch.empty_like(x)
                act = lin(m.fc_1, x, m.proj, other=m.fc_2, delta=h, x_out=x_new, norm_weight=block.norm_2.weight, norm_eps=eps)
                x = x_new
                delta = self._all_reduce(lin(m.proj, act, nxt_qkv))
            else:
                x = lin(a.proj, y, m.fc_1, residual=x)
                if split_norm:
                    act = lin(m.fc_1, K.rmsnorm(x, block.norm_2.weight, eps), m.proj, other=m.fc_2)
                else:
                    act = lin(m.fc_1, x, m.proj, other=m.fc_2, norm_weight=block.norm_2.weight, norm_eps=eps)
                x = lin(m.proj, act, nxt_qkv, residual=x)
        first = blocks[0].attn.attn if blocks else None  # graph replays run back to back: wrap the prefetch around
        if split_norm:
            logits = lin(self.lm_head, K.rmsnorm(x, self.transformer.ln_f.weight, eps), first)
        else:
            logits = lin(self.lm_head, x, first, delta=delta, norm_weight=self.transformer.ln_f.weight, norm_eps=eps)
        if tp and getattr(self, "lm_head_sharded", False):   # all-NCCL tensor-parallel path (B200_TP_NCCL=1): gather the vocabulary slices
            import torch.distributed as dist

            parts = [torch.empty_like(logits) for _ in range(self.tp_world_size)]
            dist.all_gather(parts, logits.contiguous(), group=self.tp_group)
            logits = torch.cat(parts, dim=-1)
        return logits.view(B, 1, -1)


class Block(nn.Module):
    def __init__(self, config: Config) -> None:
        super().__init__()
        self.norm_1 = config.norm_class(config.n_embd, eps=config.norm_eps)
        self.attn = CausalSelfAttention(config)
        self.norm_2 = None if config.shared_attention_norm else config.norm_class(config.n_embd, eps=config.norm_eps)
        self.mlp = config.mlp_class(config)
        self.config = config

    def forward(self, x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, mask: Optional[torch.Tensor] = None,
                input_pos: Optional[torch.Tensor] = None) -> torch.Tensor:
        n_1 = self.norm_1(x)
        h = self.attn(n_1, cos, sin, mask, input_pos)
        if self.config.parallel_residual:
            n_2 = n_1 if self.config.shared_attention_norm else self.norm_2(x)
            x = self.mlp(n_2) + h + x
        else:
            if self.config.shared_attention_norm:
                raise NotImplementedError(
                    "No checkpoint amongst the ones we support uses this configuration"
                    " (non-parallel residual and shared attention norm)."
                )
            x = h + x
            x = self.mlp(self.norm_2(x)) + x
        return x


class CausalSelfAttention(nn.Module):
    def __init__(self, config: Config) -> None:
        super().__init__()
        shape = (config.n_head + 2 * config.n_query_groups) * config.head_size
        # fused q/k/v projection; rows are interleaved per query group: [q * q_per_kv, k, v] x n_query_groups
        self.attn = nn.Linear(config.n_embd, shape, bias=config.bias)
        self.proj = nn.Linear(config.n_embd, config.n_embd, bias=config.bias)
        self.kv_cache: Optional[KVCache] = None
        self.config = config

    def forward(self, x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, mask: Optional[torch.Tensor] = None,
                input_pos: Optional[torch.Tensor] = None) -> torch.Tensor:
        B, T, C = x.size()
        c = self.config
        qkv = self.attn(x)

        if input_pos is not None:
            if not isinstance(self.kv_cache, KVCache):
                raise TypeError("You need to call `gpt.set_kv_cache()`")
            if self.kv_cache.gqa_native:
                if not (qkv.is_cuda and qkv.dtype == torch.bfloat16):
                    raise RuntimeError("the GQA-native kv cache is driven by the CUDA bf16 kernels only")
                return self.proj(self._fused_cached_attention(qkv, cos, sin, mask, input_pos))

        # ---- reference formulation (model.py:621-653) ---------------------------------------------------
        q_per_kv = c.n_head // c.n_query_groups
        total_qkv = q_per_kv + 2
        qkv = qkv.view(B, T, c.n_query_groups, total_qkv, c.head_size).permute(0, 2, 3, 1, 4)
        q, k, v = qkv.split((q_per_kv, 1, 1), dim=2)
        if c.n_query_groups != c.n_head and (input_pos is None or c.n_query_groups != 1):
            k = k.expand(B, c.n_query_groups, q_per_kv, T, c.head_size)
            v = v.expand(B, c.n_query_groups, q_per_kv, T, c.head_size)
        q = q.reshape(B, -1, T, c.head_size)
        k = k.reshape(B, -1, T, c.head_size)
        v = v.reshape(B, -1, T, c.head_size)
        q_roped = apply_rope(q[..., : c.rope_n_elem], cos, sin)
        k_roped = apply_rope(k[..., : c.rope_n_elem], cos, sin)
        q = torch.cat((q_roped, q[..., c.rope_n_elem :]), dim=-1)
        k = torch.cat((k_roped, k[..., c.rope_n_elem :]), dim=-1)
        if input_pos is not None:
            k, v = self.kv_cache(input_pos, k, v)
        y = self.scaled_dot_product_attention(q, k, v, mask)
        y = y.reshape(B, T, c.n_embd)
        return self.proj(y)

    def _fused_cached_attention(self, qkv: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor,
                                mask: Optional[torch.Tensor], input_pos: torch.Tensor) -> torch.Tensor:
        c, cache = self.config, self.kv_cache
        B, T, _ = qkv.shape
        if torch.compiler.is_compiling():  # traced: opaque b200gpt::* nodes with declared cache mutation (ops.py), no graph break
            from . import ops  # noqa: F401

            o = torch.ops.b200gpt
            if T == 1:
                y = o.attn_decode(qkv, cos, sin, input_pos, cache.k, cache.v, c.n_head, c.n_query_groups, c.head_size, c.rope_n_elem, True)
                return y.view(B, 1, c.n_head * c.head_size)
            q = o.rope_kv_append(qkv, cos, sin, input_pos, cache.k, cache.v, c.n_head, c.n_query_groups, c.head_size, c.rope_n_elem, True)
            return o.attn_prefill(q, cache.k, cache.v, input_pos, 1.0 / math.sqrt(c.head_size))
        if T == 1:
            y = K.attn_decode(qkv, cos, sin, input_pos, cache.k, cache.v, c.n_head, c.n_query_groups, c.head_size,
                              c.rope_n_elem, table_preselected=True)
            return y.view(B, 1, c.n_head * c.head_size)
        q = K.rope_kv_append(qkv, cos, sin, input_pos, cache.k, cache.v, c.n_head, c.n_query_groups, c.head_size,
                             c.rope_n_elem, table_preselected=True)
        # the mask row of token t (model.py:509) is "cache positions <= input_pos[t]": the kernel takes input_pos directly
        return K.attn_prefill(q, cache.k, cache.v, input_pos, 1.0 / math.sqrt(c.head_size))

    def scaled_dot_product_attention(self, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor,
                                     mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        scale = 1.0 / math.sqrt(self.config.head_size)
        y = F.scaled_dot_product_attention(q, k, v, attn_mask=mask, dropout_p=0.0, scale=scale, is_causal=mask is None)
        return y.transpose(1, 2)

    def build_kv_cache(self, batch_size: int, max_seq_length: int, rope_cache_length: Optional[int] = None,
                       device: Optional[torch.device] = None, dtype: Optional[torch.dtype] = None) -> "KVCache":
        c = self.config
        probe = torch.empty(0, device=device, dtype=dtype)
        native = (probe.is_cuda and probe.dtype == torch.bfloat16 and c.head_size in FUSED_HEAD_SIZES
                  and c.rope_n_elem % 2 == 0 and (rope_cache_length is None or rope_cache_length == c.rope_n_elem))
        if native:  # GQA-native: n_query_groups heads (the reference stores n_head)
            shape = (batch_size, c.n_query_groups, max_seq_length, c.head_size)
            return KVCache(shape, shape, device=device, dtype=dtype, gqa_native=True)
        heads = 1 if c.n_query_groups == 1 else c.n_head
        v_shape = (batch_size, heads, max_seq_length, c.head_size)
        if rope_cache_length is None:
            if c.rotary_percentage != 1.0:
                raise TypeError("Please pass the `rope_cache_length=gpt.cos.size(-1)` value")
            k_shape = v_shape
        else:
            k_shape = (batch_size, heads, max_seq_length, rope_cache_length + c.head_size - c.rope_n_elem)
        return KVCache(k_shape, v_shape, device=device, dtype=dtype)


class GptNeoxMLP(nn.Module):
    def __init__(self, config: Config) -> None:
        super().__init__()
        self.fc = nn.Linear(config.n_embd, config.intermediate_size, bias=config.bias)
        self.proj = nn.Linear(config.intermediate_size, config.n_embd, bias=config.bias)
        self.config = config

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.fc(x)
        x = F.gelu(x, approximate=self.config.gelu_approximate)
        return self.proj(x)


class LLaMAMLP(nn.Module):
    def __init__(self, config: Config) -> None:
        super().__init__()
        self.fc_1 = nn.Linear(config.n_embd, config.intermediate_size, bias=config.bias)
        self.fc_2 = nn.Linear(config.n_embd, config.intermediate_size, bias=config.bias)
        self.proj = nn.Linear(config.intermediate_size, config.n_embd, bias=config.bias)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        f1, f2 = self.fc_1, self.fc_2
        rows = x.numel() // x.size(-1)
        if (isinstance(f1, QLinear) and isinstance(f2, QLinear) and f1.is_quantized and f2.is_quantized and x.is_cuda
                and x.dtype == torch.bfloat16 and rows <= 16 and f1.bias is None and f2.bias is None
                and f1.quant_type == f2.quant_type and f1.quant_type != "int8" and not torch.compiler.is_compiling()):
            # gate and up projections + SwiGLU in one pass over both weight matrices
            return self.proj(f1.gemv(x, other=f2))
        x_fc_1 = self.fc_1(x)
        x_fc_2 = self.fc_2(x)
        if (x_fc_1.is_cuda and x_fc_1.dtype == torch.bfloat16 and x_fc_1.numel() % 8 == 0 and not torch.compiler.is_compiling()
                and isinstance(f1, QLinear)):
            x = K.swiglu(x_fc_1, x_fc_2)   # one kernel, the reference's rounding points (silu to bf16, product to bf16)
        else:
            x = F.silu(x_fc_1) * x_fc_2
        return self.proj(x)


class LLaMAMoE(nn.Module):
    """Mixtral-style routing is outside the hot path (no BASELINE config uses it); kept so that
    ``config._mlp_class = "LLaMAMoE"`` resolves and fails with a clear message."""

    def __init__(self, config: Config) -> None:
        super().__init__()
        raise NotImplementedError("LLaMAMoE is out of scope for the B200 quantized-decode hot path")


def build_rope_cache(seq_len: int, n_elem: int, device: Optional[torch.device] = None, base: int = 10000,
                     condense_ratio: int = 1) -> Tuple[torch.Tensor, torch.Tensor]:
    """cos/sin tables (seq_len, n_elem).  Positions are an integer arange divided by a python int, i.e. they take
    the DEFAULT float dtype exactly as the reference does (model.py:759) — the kernels consume these
    tables and never recompute angles."""
    theta = 1.0 / (base ** (torch.arange(0, n_elem, 2, device=device).float() / n_elem))
    seq_idx = torch.arange(seq_len, device=device) / condense_ratio
    idx_theta = torch.outer(seq_idx, theta).repeat(1, 2)
    return torch.cos(idx_theta), torch.sin(idx_theta)


def apply_rope(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor) -> torch.Tensor:
    head_size = x.size(-1)
    x1 = x[..., : head_size // 2]
    x2 = x[..., head_size // 2 :]
    rotated = torch.cat((-x2, x1), dim=-1)
    roped = (x * cos) + (rotated * sin)
    return roped.to(dtype=x.dtype)


class KVCache(nn.Module):
    def __init__(self, k_shape: Tuple[int, int, int, int], v_shape: Tuple[int, int, int, int],
                 device: Optional[torch.device] = None, dtype: Optional[torch.dtype] = None,
                 gqa_native: bool = False) -> None:
        super().__init__()
        self.register_buffer("k", torch.zeros(k_shape, device=device, dtype=dtype), persistent=False)
        self.register_buffer("v", torch.zeros(v_shape, device=device, dtype=dtype), persistent=False)
        self.gqa_native = gqa_native

    def forward(self, input_pos: torch.Tensor, k: torch.Tensor, v: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        if self.gqa_native:
            raise RuntimeError("a GQA-native cache is updated inside the fused attention kernels")
        self.k = self.k.to(k.dtype)
        self.v = self.v.to(v.dtype)
        k = self.k.index_copy_(2, input_pos, k)
        v = self.v.index_copy_(2, input_pos, v)
        return k, v

    def reset_parameters(self) -> None:
        torch.nn.init.zeros_(self.k)
        torch.nn.init.zeros_(self.v)


def build_mask_cache(max_seq_length: int, device: Optional[torch.device] = None) -> torch.Tensor:
    ones = torch.ones((max_seq_length, max_seq_length), device=device, dtype=torch.bool)
    return torch.tril(ones).unsqueeze(0).unsqueeze(0)
