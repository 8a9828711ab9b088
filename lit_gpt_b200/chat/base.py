"""Streaming generator over ``next_token`` -- the second caller of the decode hot path (reference chat/base.py:24-68).

Same contract as the reference's ``chat.base.generate``: yields the new tokens one by one, holds back as many tokens as the
longest stop sequence so that a stop sequence is never emitted, and returns as soon as one matches.  On the fused quantized
path every step is one replay of the captured decode graph (``generate.base.DecodeGraph``); the only host synchronisation is
the read of the new token that a streaming caller needs anyway.  The prompt / UI / tokenizer parts of chat/base.py are out of
scope.
"""
from __future__ import annotations

from typing import Iterator, List, Optional, Tuple

import torch

import lit_gpt_b200.generate.base as generate_base
from lit_gpt_b200 import GPT


@torch.inference_mode()
def generate(
    model: GPT,
    prompt: torch.Tensor,
    max_returned_tokens: int,
    *,
    temperature: float = 1.0,
    top_k: Optional[int] = None,
    stop_tokens: Tuple[List[int], ...] = (),
) -> Iterator[torch.Tensor]:
    T = prompt.size(0)
    assert max_returned_tokens > T
    if model.max_seq_length < max_returned_tokens - 1:
        raise NotImplementedError(f"max_seq_length {model.max_seq_length} needs to be >= {max_returned_tokens - 1}")

    device = prompt.device
    buffer_length = max((len(tokens) for tokens in stop_tokens), default=1)
    stops = [list(st) for st in stop_tokens]
    yield_i = 0
    tokens: List[torch.Tensor] = []
    values: List[int] = []          # host copies of the generated ids: the stop test runs on the host

    def stopped() -> bool:
        return any(len(st) <= len(values) and values[-len(st):] == st for st in stops)

    # prefill through next_token (chat/base.py:55-58), then one decode step per token
    input_pos = torch.arange(0, T, device=device)
    token = generate_base.next_token(model, input_pos, prompt.view(1, -1), temperature=temperature, top_k=top_k)
    dg = None
    fused = getattr(model, "fused_decode_ready", None)
    if prompt.is_cuda and fused is not None and model.fused_decode_ready(prompt.view(1, -1)) and max_returned_tokens - T > 1 \
            and not model.spans_devices():
        dg = generate_base._decode_graph_for(model, temperature, top_k, prompt.dtype)
        dg.token.copy_(token.view(-1))
        dg.pos.fill_(T)
        if dg.graph is None:
            dg.capture()
    input_pos = input_pos[-1:].add_(1)
    for t in range(1, max_returned_tokens - T + 1):
        if t > 1:
            if dg is not None:
                dg.replay()
                token = dg.token.clone()
            else:
                token = generate_base.next_token(model, input_pos, token.view(1, -1), temperature=temperature, top_k=top_k)
                input_pos = input_pos.add_(1)
        tokens.append(token)
        if stops:
            values.append(int(token))
            if stopped():
                return
        if t - yield_i >= buffer_length:
            yield from tokens[yield_i:t]
            yield_i = t
