"""Model hyper-parameters: same public surface as the reference's ``lit_gpt/config.py:16-144`` (field names,
``from_name`` / ``from_json`` / ``from_checkpoint``, derived ``head_size`` / ``rope_n_elem`` /
``padded_vocab_size``) so a ``lit_config.json`` written by the reference loads unchanged.

The registry below holds the model families on the hot path named by BASELINE.json (pythia for the plumbing config,
Llama-2 / CodeLlama / Mistral / TinyLlama for the quantized decode configs); other families can be passed as kwargs
or through ``lit_config.json``.
"""
from __future__ import annotations

import json
from dataclasses import dataclass, field
from pathlib import Path
from typing import Any, Dict, List, Literal, Optional, Type, Union

import torch


def find_multiple(n: int, k: int) -> int:
    """Smallest multiple of k that is >= n (reference lit_gpt/utils.py find_multiple)."""
    return n if n % k == 0 else n + k - (n % k)


@dataclass
class Config:
    name: str = ""
    hf_config: dict = field(default_factory=dict)
    block_size: int = 4096
    vocab_size: int = 50254
    padding_multiple: int = 512
    padded_vocab_size: Optional[int] = None
    n_layer: int = 16
    n_head: int = 32
    n_embd: int = 4096
    rotary_percentage: float = 0.25
    parallel_residual: bool = True
    bias: bool = True
    lm_head_bias: bool = False
    # n_query_groups == n_head: MHA; 1: MQA; in between: GQA (reference config.py:31-51)
    n_query_groups: Optional[int] = None
    shared_attention_norm: bool = False
    _norm_class: Literal["LayerNorm", "RMSNorm"] = "LayerNorm"
    norm_eps: float = 1e-5
    _mlp_class: Literal["GptNeoxMLP", "LLaMAMLP"] = "GptNeoxMLP"
    gelu_approximate: str = "none"
    intermediate_size: Optional[int] = None
    rope_condense_ratio: int = 1
    rope_base: int = 10000
    n_expert: int = 0
    n_expert_per_token: int = 0

    def __post_init__(self) -> None:
        if not self.name:
            self.name = self.hf_config.get("name", self.name)
        assert self.n_embd % self.n_head == 0
        self.head_size = self.n_embd // self.n_head
        if self.padded_vocab_size is None:
            self.padded_vocab_size = find_multiple(self.vocab_size, self.padding_multiple)
        else:
            self.vocab_size = min(self.vocab_size, self.padded_vocab_size)
        if self.n_query_groups is not None:
            assert self.n_head % self.n_query_groups == 0
        else:
            self.n_query_groups = self.n_head
        if self.intermediate_size is None:
            if self._mlp_class == "LLaMAMLP":
                raise ValueError("The config needs to set the `intermediate_size`")
            self.intermediate_size = 4 * self.n_embd
        self.rope_n_elem = int(self.rotary_percentage * self.head_size)

    @classmethod
    def from_name(cls, name: str, **kwargs: Any) -> "Config":
        if name in name_to_config:
            conf = dict(name_to_config[name])
        else:
            matches = [c for c in configs if c["hf_config"]["name"] == name]
            if not matches:
                raise ValueError(f"{name!r} is not a supported config name")
            conf = dict(matches[0])
        if "condense_ratio" in kwargs:  # legacy key
            kwargs["rope_condense_ratio"] = kwargs.pop("condense_ratio")
        conf.update(kwargs)
        return cls(**conf)

    @classmethod
    def from_json(cls, path: Union[str, Path], **kwargs: Any) -> "Config":
        with open(path, encoding="utf-8") as fp:
            js = json.load(fp)
        for d in (js, kwargs):
            if "condense_ratio" in d:
                d["rope_condense_ratio"] = d.pop("condense_ratio")
        if "org" in js:
            js["hf_config"] = {"name": js["name"], "org": js.pop("org")}
        if "org" in kwargs:
            kwargs["hf_config"] = {"name": kwargs.get("name", js["name"]), "org": kwargs.pop("org")}
        js.update(kwargs)
        return cls(**js)

    @classmethod
    def from_checkpoint(cls, path: Path, **kwargs: Any) -> "Config":
        if (config_path := path / "lit_config.json").is_file():
            return cls.from_json(config_path, **kwargs)
        if (model_name := path.name) in name_to_config:
            return cls.from_name(model_name, **kwargs)
        raise FileNotFoundError(f"For {str(path)!r} neither 'lit_config.json' nor matching config exists.")

    @property
    def mlp_class(self) -> Type:
        import lit_gpt_b200.model as model

        return getattr(model, self._mlp_class)

    @property
    def norm_class(self) -> Type:
        if self._norm_class == "RMSNorm":
            from lit_gpt_b200.rmsnorm import RMSNorm

            return RMSNorm
        return getattr(torch.nn, self._norm_class)


# ---------------------------------------------------------------------------------------------------------
# registry
# ---------------------------------------------------------------------------------------------------------
configs: List[Dict[str, Any]] = []


def _add(name: str, org: str, hf_name: Optional[str] = None, **kw: Any) -> None:
    configs.append(dict(name=name, hf_config=dict(org=org, name=hf_name or name), **kw))


# EleutherAI Pythia (GPT-NeoX: LayerNorm, GELU MLP, bias, parallel residual, 25 % rotary)
_PYTHIA = {
    "pythia-14m": dict(block_size=512, n_layer=6, n_embd=128, n_head=4, padding_multiple=128),
    "pythia-31m": dict(block_size=1024, n_layer=6, n_embd=256, n_head=8, padding_multiple=128),
    "pythia-70m": dict(block_size=2048, n_layer=6, n_embd=512, n_head=8, padding_multiple=128),
    "pythia-160m": dict(block_size=2048, n_layer=12, n_embd=768, n_head=12, padding_multiple=128),
    "pythia-410m": dict(block_size=2048, n_layer=24, n_embd=1024, n_head=16, padding_multiple=128),
    "pythia-1b": dict(block_size=2048, n_embd=2048, n_head=8, padding_multiple=128),
    "pythia-1.4b": dict(block_size=2048, n_layer=24, n_embd=2048, n_head=16, padding_multiple=128),
    "pythia-2.8b": dict(block_size=2048, n_layer=32, n_embd=2560, padding_multiple=128),
    "pythia-6.9b": dict(block_size=2048, n_layer=32, padding_multiple=256),
    "pythia-12b": dict(block_size=2048, n_layer=36, n_embd=5120, n_head=40),
}
for _n, _kw in _PYTHIA.items():
    _add(_n, "EleutherAI", **_kw)
    if _n not in ("pythia-14m", "pythia-31m"):
        _add(f"{_n}-deduped", "EleutherAI", **_kw)

# Llama-family decoder: RMSNorm, SwiGLU MLP, no bias, sequential residual, full rotary
_LLAMA = dict(rotary_percentage=1.0, parallel_residual=False, bias=False, _norm_class="RMSNorm", _mlp_class="LLaMAMLP")
_LLAMA2 = {
    "7b": dict(n_layer=32, intermediate_size=11008),
    "13b": dict(n_layer=40, n_head=40, n_embd=5120, intermediate_size=13824),
    "70b": dict(n_layer=80, n_head=64, n_embd=8192, n_query_groups=8, intermediate_size=28672),
}
for _size, _kw in _LLAMA2.items():
    for _kind in ("", "-chat"):
        _add(f"Llama-2-{_size}{_kind}-hf", "meta-llama", vocab_size=32000, padding_multiple=64, **_LLAMA, **_kw)

_CODELLAMA = {
    "7b": dict(n_layer=32, intermediate_size=11008),
    "13b": dict(n_layer=40, n_head=40, n_embd=5120, intermediate_size=13824),
    "34b": dict(n_layer=48, n_head=64, n_embd=8192, n_query_groups=8, intermediate_size=22016),
}
for _size, _kw in _CODELLAMA.items():
    for _kind, _vocab in (("", 32016), ("-Python", 32000), ("-Instruct", 32016)):
        _v = 32000 if _size == "34b" else _vocab
        _add(f"CodeLlama-{_size}{_kind}-hf", "codellama", vocab_size=_v, padding_multiple=16, block_size=16384,
             rope_base=1000000, **_LLAMA, **_kw)

for _kind in ("", "Instruct-"):
    # block_size 4096: sliding-window attention is not implemented (same limit as the reference, config.py:1278)
    _add(f"Mistral-7B-{_kind}v0.1", "mistralai", padded_vocab_size=32000, block_size=4096, n_layer=32, n_query_groups=8,
         norm_eps=1e-05, intermediate_size=14336, **_LLAMA)

for _kind, _hf in (("", "-intermediate-step-1431k-3T"), ("-chat", "-Chat-v1.0")):
    _add(f"tiny-llama-1.1b{_kind}", "TinyLlama", hf_name=f"TinyLlama-1.1B{_hf}", block_size=2048, vocab_size=32000,
         padding_multiple=64, n_layer=22, n_head=32, n_embd=2048, n_query_groups=4, intermediate_size=5632, **_LLAMA)

name_to_config: Dict[str, Dict[str, Any]] = {c["name"]: c for c in configs}
