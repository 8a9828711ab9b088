"""``generate/base.py`` of the reference (lines 22-194) on the B200 path: same ``generate`` / ``next_token`` /
``sample`` / ``main`` signatures, the same ``--quantize`` / ``--precision`` flags and the same stderr timing lines.

When the model is on the fused quantized path, the per-token loop (reference base.py:87-92: one python iteration,
~45 kernel launches per block and one device->host sync per token) becomes replays of ONE captured CUDA graph whose
nodes are this library's kernels chained by programmatic dependent launch; the sampled token, ``input_pos`` and the
output sequence all stay on the device, and the EOS test is polled every ``EOS_POLL`` steps instead of every step.
"""
from __future__ import annotations

import sys
import time
from pathlib import Path
from typing import Any, Dict, Literal, Optional, Tuple

import torch

from lit_gpt_b200 import GPT, Config, QuantPrecision
from lit_gpt_b200 import kernels as K
from lit_gpt_b200.fabric import Fabric, seed_everything
from lit_gpt_b200.quant import weight_epoch
from lit_gpt_b200.utils import check_valid_checkpoint_dir, get_default_supported_precision, load_checkpoint

EOS_POLL = 16


def multinomial_num_samples_1(probs: torch.Tensor) -> torch.Tensor:
    if (probs.is_cuda and torch.cuda.is_current_stream_capturing()) or torch.compiler.is_compiling():
        # CUDA-graph friendly draw (the reference uses the same trick under torch.compile, base.py:23-26)
        distribution = torch.empty_like(probs).exponential_(1)
        return torch.argmax(probs / distribution, dim=-1, keepdim=True)
    return torch.multinomial(probs, num_samples=1)


_RNG: Dict[torch.device, Tuple[torch.Tensor, int, int]] = {}


def _rng_for(device: torch.device) -> torch.Tensor:
    """ONE Philox state per device, shared by the prefill draw and every captured decode graph (separate states with the same seed
    would replay the same noise: correlated samples).  Reseeded IN PLACE -- captured graphs hold its address -- whenever torch's
    seed changes or ``seed_everything`` is called again, so that a re-seeded run reproduces like the reference's."""
    import lit_gpt_b200.fabric as _fabric

    device = torch.device(device.type, device.index if device.index is not None else torch.cuda.current_device())
    seed, epoch = torch.initial_seed() & 0x7FFFFFFFFFFFFFFF, _fabric.SEED_EPOCH
    hit = _RNG.get(device)
    if hit is None:
        st = K.new_rng_state(device)
    elif hit[1] != seed or hit[2] != epoch:
        st = hit[0]
        st.copy_(torch.tensor([seed, 0], dtype=torch.int64))
    else:
        return hit[0]
    _RNG[device] = (st, seed, epoch)
    return st


def sample(logits: torch.Tensor, temperature: float = 1.0, top_k: Optional[int] = None) -> torch.Tensor:
    if logits.is_cuda and logits.dtype in (torch.bfloat16, torch.float32) and logits.dim() == 3:
        # one fused launch (csrc/sample.cu): top-k threshold, temperature softmax, the Exp(1) draw of base.py:23-26
        out = torch.empty(1, dtype=torch.int64, device=logits.device)
        K.sample_(logits, out, temperature, top_k, None if temperature <= 0.0 else _rng_for(logits.device))
        return out
    logits = logits[0, -1]
    if top_k is not None:
        v, i = torch.topk(logits, min(top_k, logits.size(-1)))
        # scatter, not `where`: repeated top-k values must not be duplicated
        logits = torch.full_like(logits, float("-inf")).scatter_(-1, i, v)
    if temperature > 0.0:
        probs = torch.nn.functional.softmax(logits / temperature, dim=-1)
        return multinomial_num_samples_1(probs)
    return torch.argmax(logits, dim=-1, keepdim=True)


def next_token(model: GPT, input_pos: torch.Tensor, x: torch.Tensor, **kwargs: Any) -> torch.Tensor:
    logits = model(x, input_pos)
    next = sample(logits, **kwargs)
    return next.to(dtype=x.dtype)


class DecodeGraph:
    """One captured decode step: token/pos/history live on the device; ``replay()`` advances one token."""

    def __init__(self, model: GPT, temperature: float, top_k: Optional[int], dtype: torch.dtype) -> None:
        dev = model.transformer.wte.weight.device
        self.model, self.temperature, self.top_k = model, temperature, top_k
        self.token = torch.zeros(1, dtype=dtype, device=dev)
        self.pos = torch.zeros(1, dtype=torch.int64, device=dev)
        self.history = torch.zeros(model.max_seq_length + 1, dtype=dtype, device=dev)
        self.max_seq_length = model.max_seq_length
        self.rng = _rng_for(dev)   # the device's one sampler state (shared with the prefill draw)
        self.graph: Optional[torch.cuda.CUDAGraph] = None

    def _step(self) -> None:
        logits = self.model(self.token.view(1, 1), self.pos)
        if self.temperature == 0.0 and self.top_k is None:
            K.argmax_(logits, self.token, pos_inout=self.pos, history=self.history)  # greedy: fused argmax + bookkeeping
        else:  # stochastic: the fused sampler draws, advances its Philox offset and does the same bookkeeping
            K.sample_(logits, self.token, self.temperature, self.top_k, self.rng, pos_inout=self.pos, history=self.history)

    def capture(self) -> None:
        state = (self.token.clone(), self.pos.clone())
        rng0 = self.rng.clone()   # the warm-up steps below draw from the sampler state: put it back, they are not part of the sequence
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):  # warm-up outside capture: allocates workspaces, sets kernel attributes
            for _ in range(2):
                self.token.copy_(state[0]); self.pos.copy_(state[1])
                self._step()
        torch.cuda.current_stream().wait_stream(side)
        self.token.copy_(state[0]); self.pos.copy_(state[1])
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self._step()
        self.token.copy_(state[0]); self.pos.copy_(state[1])
        self.rng.copy_(rng0)

    def replay(self) -> None:
        self.graph.replay()


PREFILL_GRAPH_MAX_T = 16   # prompts up to this length run their prefill as one captured graph (0 switches it off)


class PrefillGraph:
    """The prefill of a SHORT prompt (T <= 16 rows: every Linear is a batched GEMV launch) plus the first draw, captured once per
    prompt length.  Run eagerly that prefill is host-bound: ~300 launches at ~17 us of Python / ctypes each (5 - 6 ms for a 7B
    model) against ~2 ms on the device; `generate/base.py`'s sample loop and a served short prompt repeat the same length.  The
    first call with a new length pays the capture (two warm-up runs + the capture, ~3 x an eager prefill)."""

    def __init__(self, model: GPT, T: int, temperature: float, top_k: Optional[int], dtype: torch.dtype) -> None:
        dev = model.transformer.wte.weight.device
        self.model, self.temperature, self.top_k = model, temperature, top_k
        self.prompt = torch.zeros(1, T, dtype=dtype, device=dev)
        self.pos = torch.arange(0, T, device=dev)
        self.token: Optional[torch.Tensor] = None
        self.rng = _rng_for(dev)
        self.graph: Optional[torch.cuda.CUDAGraph] = None

    def _run(self) -> None:
        self.token = next_token(self.model, self.pos, self.prompt, temperature=self.temperature, top_k=self.top_k)

    def capture(self) -> None:
        rng0 = self.rng.clone()   # the warm-up runs draw from the sampler state: put it back (DecodeGraph.capture)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(2):
                self._run()
        torch.cuda.current_stream().wait_stream(side)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self._run()
        self.rng.copy_(rng0)


def _graph_key(model: GPT) -> Tuple:
    # a captured graph bakes in raw device pointers: key on the addresses themselves (id() of a freed tensor can be reused)
    kv = [(b.attn.kv_cache.k.data_ptr(), b.attn.kv_cache.v.data_ptr()) for b in model.transformer.h]
    dev = model.transformer.wte.weight.device
    rng = _rng_for(dev).data_ptr() if dev.type == "cuda" else 0   # the sampler state's address is baked in too
    return (model.max_seq_length, tuple(kv), model.cos.data_ptr(), model.sin.data_ptr(), weight_epoch(), rng)


def _prefill_graph_for(model: GPT, T: int, temperature: float, top_k: Optional[int], dtype: torch.dtype) -> PrefillGraph:
    cache: Dict[Tuple, PrefillGraph] = model.__dict__.setdefault("_prefill_graphs", {})
    key = (T, temperature, top_k, dtype) + _graph_key(model)
    pg = cache.get(key)
    if pg is None:
        if cache and next(iter(cache))[4:] != key[4:]:
            cache.clear()  # buffers or weights changed: the old graphs pin stale addresses
        if len(cache) >= 2 * PREFILL_GRAPH_MAX_T:
            cache.pop(next(iter(cache)))
        pg = PrefillGraph(model, T, temperature, top_k, dtype)
        cache[key] = pg
    return pg


def _decode_graph_for(model: GPT, temperature: float, top_k: Optional[int], dtype: torch.dtype) -> DecodeGraph:
    cache: Dict[Tuple, DecodeGraph] = model.__dict__.setdefault("_decode_graphs", {})
    key = (temperature, top_k, dtype) + _graph_key(model)
    dg = cache.get(key)
    if dg is None:
        cache.clear()  # graphs pin the kv-cache buffers they were captured with
        dg = DecodeGraph(model, temperature, top_k, dtype)
        cache[key] = dg
    return dg


@torch.inference_mode()
def generate(model: GPT, prompt: torch.Tensor, max_returned_tokens: int, *, temperature: float = 1.0,
             top_k: Optional[int] = None, eos_id: Optional[int] = None) -> torch.Tensor:
    """prompt: (T,) token ids.  Returns prompt + generated ids, at most ``max_returned_tokens`` long; stops after
    ``eos_id``.  Same contract as the reference (generate/base.py:51-93)."""
    T = prompt.size(0)
    assert max_returned_tokens > T
    if model.max_seq_length < max_returned_tokens - 1:
        # a rolling cache is not implemented (same limit and message as the reference, base.py:74-78)
        raise NotImplementedError(f"max_seq_length {model.max_seq_length} needs to be >= {max_returned_tokens - 1}")

    device = prompt.device
    if device.type == "cuda" and temperature > 0.0:
        _rng_for(device)   # apply a pending reseed before any draw of this call
    tokens = [prompt]
    input_pos = torch.tensor([T], device=device)
    if (1 < T <= PREFILL_GRAPH_MAX_T and prompt.is_cuda and getattr(model, "tp_world_size", 1) == 1 and not model.spans_devices()
            and model.fused_decode_ready(prompt.view(1, -1)) and model.config.intermediate_size % 8 == 0):
        pg = _prefill_graph_for(model, T, temperature, top_k, prompt.dtype)
        pg.prompt.copy_(prompt.view(1, -1))
        if pg.graph is None:
            pg.capture()
        pg.graph.replay()
        token = pg.token.clone()
    else:
        token = next_token(
            model, torch.arange(0, T, device=device), prompt.view(1, -1), temperature=temperature, top_k=top_k
        ).clone()
    tokens.append(token)
    n_steps = max_returned_tokens - T - 1

    # (a model whose blocks sit on several devices -- generate/sequentially.py -- is driven by one process: no single CUDA graph)
    if n_steps > 0 and prompt.is_cuda and model.fused_decode_ready(prompt.view(1, -1)) and not model.spans_devices():
        dg = _decode_graph_for(model, temperature, top_k, prompt.dtype)
        dg.token.copy_(token)
        dg.pos.fill_(T)
        if dg.graph is None:
            dg.capture()
        done = 0
        n_valid = n_steps
        while done < n_steps:
            burst = n_steps - done if eos_id is None else min(EOS_POLL, n_steps - done)
            for _ in range(burst):
                dg.replay()
            done += burst
            if eos_id is not None:  # the only host sync of the loop, every EOS_POLL tokens
                seg = dg.history[T + 1 : T + 1 + done]
                hit = (seg == eos_id).nonzero()
                if hit.numel():
                    n_valid = int(hit[0]) + 1
                    break
        if n_valid:
            tokens.append(dg.history[T + 1 : T + 1 + min(n_valid, done)].clone())
        if getattr(model, "tp_world_size", 1) > 1:  # a peer that died mid-sequence shows up here, not as a hung GPU
            K.check_gemv_errors(device)
        return torch.cat(tokens)

    for _ in range(2, max_returned_tokens - T + 1):
        token = next_token(model, input_pos, token.view(1, -1), temperature=temperature, top_k=top_k).clone()
        tokens.append(token)
        if token == eos_id:
            break
        input_pos = input_pos.add_(1)
    return torch.cat(tokens)


@torch.inference_mode()
def main(
    prompt: str = "What food do llamas eat?",
    *,
    num_samples: int = 1,
    max_new_tokens: int = 50,
    top_k: Optional[int] = 200,
    temperature: float = 0.8,
    checkpoint_dir: Path = Path("checkpoints/stabilityai/stablelm-base-alpha-3b"),
    quantize: Optional[Literal["bnb.nf4", "bnb.nf4-dq", "bnb.fp4", "bnb.fp4-dq", "bnb.int8", "gptq.int4"]] = None,
    precision: Optional[str] = None,
    compile: bool = False,
) -> None:
    """Generates text samples based on a pre-trained model and tokenizer.

    Args mirror the reference CLI (generate/base.py:97-124).  ``quantize``: bnb.nf4 / bnb.fp4 are the bitsandbytes
    block-64 4-bit formats, gptq.int4 is this project's int4 group-128 format.  ``compile`` is accepted for parity;
    the decode loop is a CUDA graph of hand-written kernels, there is no tracing compiler on this path.
    """
    from lit_gpt_b200.tokenizer import Tokenizer

    precision = precision or get_default_supported_precision(training=False)

    plugins = None
    if quantize is not None:
        if "mixed" in precision:
            raise ValueError("Quantization and mixed precision is not supported.")
        dtype = {"16-true": torch.float16, "bf16-true": torch.bfloat16, "32-true": torch.float32}[precision]
        plugins = QuantPrecision(quantize, dtype)
        precision = None

    fabric = Fabric(devices=1, precision=precision, plugins=plugins)

    check_valid_checkpoint_dir(checkpoint_dir)
    config = Config.from_json(checkpoint_dir / "lit_config.json")
    checkpoint_path = checkpoint_dir / "lit_model.pth"

    tokenizer = Tokenizer(checkpoint_dir)
    encoded = tokenizer.encode(prompt, device=fabric.device)
    prompt_length = encoded.size(0)
    max_returned_tokens = prompt_length + max_new_tokens

    fabric.print(f"Loading model {str(checkpoint_path)!r} with {config.__dict__}", file=sys.stderr)
    t0 = time.perf_counter()
    with fabric.init_module(empty_init=True):
        model = GPT(config)
    fabric.print(f"Time to instantiate model: {time.perf_counter() - t0:.02f} seconds.", file=sys.stderr)
    with fabric.init_tensor():
        model.max_seq_length = max_returned_tokens
        model.set_kv_cache(batch_size=1)
    model.eval()

    model = fabric.setup_module(model)

    t0 = time.perf_counter()
    load_checkpoint(fabric, model, checkpoint_path)
    fabric.print(f"Time to load the model weights: {time.perf_counter() - t0:.02f} seconds.", file=sys.stderr)

    seed_everything(1234)
    for i in range(num_samples):
        t0 = time.perf_counter()
        y = generate(model, encoded, max_returned_tokens, temperature=temperature, top_k=top_k, eos_id=tokenizer.eos_id)
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


def _cli(fn) -> None:
    try:
        from jsonargparse import CLI

        CLI(fn)
    except ImportError:  # same flags through argparse when jsonargparse is absent
        import argparse
        import inspect

        ap = argparse.ArgumentParser(description=(fn.__doc__ or "").strip().splitlines()[0])
        for name, p in inspect.signature(fn).parameters.items():
            default = p.default
            kind = type(default) if default is not None and not isinstance(default, bool) else str
            if isinstance(default, bool):
                ap.add_argument(f"--{name}", type=lambda s: s.lower() in ("1", "true", "yes"), default=default)
            elif name == "prompt":
                ap.add_argument("prompt", nargs="?", default=default)
            else:
                ap.add_argument(f"--{name}", type=(Path if isinstance(default, Path) else (int if name in ("top_k", "num_samples", "max_new_tokens") else kind)), default=default)
        fn(**vars(ap.parse_args()))


if __name__ == "__main__":
    torch.set_float32_matmul_precision("high")
    _cli(main)
