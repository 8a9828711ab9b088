"""Host logic of the GPT / Config / generate mirror on CPU (fp32, unquantized): results identical to the reference's.
Mirrors the reference's tests/test_model.py, tests/test_generate.py (cited per test)."""
import ast
from unittest import mock

import numpy as np
import pytest
import torch

import lit_gpt_b200.generate.base as generate
from lit_gpt_b200 import GPT, Config
from oracle import model_ref as M

CASES = ["llama2_toy", "llama2_70b_mqa_toy", "mistral_toy", "llama_gqa_hs64", "llama_mha_hs128", "pythia_toy"]


def build(golden, name, dtype=torch.float32):
    g = golden(name)
    cfgd = ast.literal_eval(str(g["config"]))
    cfg = Config(**{k: v for k, v in cfgd.items() if k not in ("head_size", "rope_n_elem")},
                 rotary_percentage=cfgd["rope_n_elem"] / cfgd["head_size"])
    rc = M.RefConfig.from_any(cfg)
    p = M.init_weights(rc, seed=1234, bias=cfgd["bias"])
    if cfgd["_norm_class"] == "LayerNorm":
        gen = torch.Generator().manual_seed(7)
        for k in p:
            if "norm" in k or "ln_f" in k:
                p[k] = p[k] + 0.1 * torch.randn(p[k].shape, generator=gen)
    model = GPT(cfg)
    model.load_state_dict(p, strict=True)
    return g, cfg, model.eval()


@pytest.mark.parametrize("name", CASES)
@torch.inference_mode()
def test_forward_equals_reference_golden(golden, name):
    g, cfg, model = build(golden, name)
    y = model(torch.from_numpy(g["fwd_input"]))
    torch.testing.assert_close(y, torch.from_numpy(g["fwd_logits"]))


@pytest.mark.parametrize("name", CASES)
@torch.inference_mode()
def test_generate_equals_reference_golden(golden, name):
    g, cfg, model = build(golden, name)
    n = len(g["gen_tokens"])
    model.max_seq_length = n
    model.set_kv_cache(batch_size=1)
    out = generate.generate(model, torch.from_numpy(g["gen_prompt"]), n, temperature=0.0)
    assert out.tolist() == g["gen_tokens"].tolist()
    model.set_kv_cache(batch_size=2)
    x = torch.from_numpy(g["fwd_input"])
    xb = torch.stack([x[0], x[0].flip(0)])
    torch.testing.assert_close(model(xb, torch.arange(5)), torch.from_numpy(g["b2_prefill_logits"]))
    torch.testing.assert_close(model(torch.tensor([[3], [7]]), torch.tensor([5])), torch.from_numpy(g["b2_decode_logits"]))


@torch.inference_mode()
@pytest.mark.parametrize("ours_kwargs", [{"name": "Llama-2-7b-hf"}, {"name": "CodeLlama-7b-hf"},
                                         {"name": "Llama-2-70b-chat-hf", "n_query_groups": 1}])
def test_against_hf_llama2(ours_kwargs):
    """reference tests/test_model.py:201-261 with our GPT; the QKV interleave of scripts/convert_hf_checkpoint.py:177-188
    is restated inline."""
    from transformers.models.llama.configuration_llama import LlamaConfig
    from transformers.models.llama.modeling_llama import LlamaForCausalLM

    cfg = Config.from_name(padded_vocab_size=10000, n_layer=2, n_head=8, n_embd=32, intermediate_size=86, **ours_kwargs)
    T = 5
    theirs_config = LlamaConfig(
        vocab_size=cfg.padded_vocab_size, hidden_size=cfg.n_embd, num_attention_heads=cfg.n_head,
        num_hidden_layers=cfg.n_layer, intermediate_size=cfg.intermediate_size, max_position_embeddings=T,
        rms_norm_eps=cfg.norm_eps, num_key_value_heads=cfg.n_query_groups, rope_theta=cfg.rope_base,
        attention_bias=cfg.bias)
    theirs = LlamaForCausalLM(theirs_config).eval()
    sd = theirs.state_dict()
    ours = {"transformer.wte.weight": sd["model.embed_tokens.weight"], "transformer.ln_f.weight": sd["model.norm.weight"],
            "lm_head.weight": sd["lm_head.weight"]}
    q_per_kv = cfg.n_head // cfg.n_query_groups
    for i in range(cfg.n_layer):
        h, t = f"model.layers.{i}.", f"transformer.h.{i}."
        q = sd[h + "self_attn.q_proj.weight"].split(cfg.head_size * q_per_kv)
        k = sd[h + "self_attn.k_proj.weight"].split(cfg.head_size)
        v = sd[h + "self_attn.v_proj.weight"].split(cfg.head_size)
        ours[t + "attn.attn.weight"] = torch.cat([w for tpl in zip(q, k, v) for w in tpl])
        ours[t + "attn.proj.weight"] = sd[h + "self_attn.o_proj.weight"]
        ours[t + "norm_1.weight"] = sd[h + "input_layernorm.weight"]
        ours[t + "norm_2.weight"] = sd[h + "post_attention_layernorm.weight"]
        ours[t + "mlp.fc_1.weight"] = sd[h + "mlp.gate_proj.weight"]
        ours[t + "mlp.fc_2.weight"] = sd[h + "mlp.up_proj.weight"]
        ours[t + "mlp.proj.weight"] = sd[h + "mlp.down_proj.weight"]
    model = GPT(cfg).eval()
    model.load_state_dict(ours)
    x = torch.tensor([[9856, 23, 491, 1536, 304]], dtype=torch.int32)
    torch.testing.assert_close(model(x), theirs(x)["logits"])


@torch.inference_mode()
@pytest.mark.parametrize("max_seq_length", (25, pytest.param(23, marks=pytest.mark.xfail(raises=IndexError, strict=True))))
def test_kv_cache(max_seq_length):
    """reference tests/test_model.py:578-612: cached decode == full recompute, 20 greedy steps, atol 0."""
    torch.manual_seed(0)
    config = Config(block_size=25, padded_vocab_size=5, n_layer=2, n_head=2, n_embd=8)
    model = GPT(config)
    idx = torch.randint(0, model.config.padded_vocab_size, (1, 5))
    model.max_seq_length = max_seq_length
    model.set_kv_cache(1)

    def pick(logits):
        return torch.argmax(torch.softmax(logits[:, -1:], dim=-1)).unsqueeze(0).unsqueeze(0)

    x_no_cache = x_cache = idx
    input_pos = torch.arange(0, 5)
    for _ in range(20):
        out_no_cache = pick(model(x_no_cache[:, -max_seq_length:]))
        out_cache = pick(model(x_cache, input_pos))
        torch.testing.assert_close(out_no_cache, out_cache, rtol=0, atol=0)
        x_no_cache = torch.cat((x_no_cache, out_no_cache), dim=1)
        x_cache = out_cache
        input_pos = input_pos[-1:] + 1


@pytest.mark.parametrize("max_seq_length", (pytest.param(10, marks=pytest.mark.xfail(raises=NotImplementedError, strict=True)), 25))
def test_generate(max_seq_length):
    """reference tests/test_generate.py:16-46"""
    T = 5
    input_idx = torch.randint(10, size=(T,))
    model = GPT(Config(block_size=128, vocab_size=16, n_layer=1, n_head=4, n_embd=8))
    model.max_seq_length = max_seq_length
    model.set_kv_cache(batch_size=1)
    draws = []

    def multinomial(*args, **kwargs):
        out = torch.multinomial(*args, **kwargs, num_samples=1)
        draws.append(out)
        return out

    with mock.patch("lit_gpt_b200.generate.base.multinomial_num_samples_1", multinomial):
        out = generate.generate(model, input_idx, T + 20, top_k=4)
    assert out.size(0) == T + 20
    torch.testing.assert_close(out, torch.cat((input_idx, torch.hstack(draws))))


@pytest.mark.parametrize("temperature", (0.0, 1.0, 0.5))
def test_sample(temperature, golden):
    """reference tests/test_generate.py:94-107"""
    logits = torch.from_numpy(golden("units")["sample_logits"])
    token = generate.sample(logits, temperature=temperature)
    assert token.shape == (1,) and token.tolist() == [0]


def test_errors_match_reference():
    model = GPT(Config(block_size=16, vocab_size=16, n_layer=1, n_head=2, n_embd=8))
    with pytest.raises(TypeError, match="set_kv_cache"):
        model(torch.zeros(1, 1, dtype=torch.long), torch.tensor([0]))
    with pytest.raises(ValueError, match="Cannot forward sequence of length"):
        model(torch.zeros(1, 17, dtype=torch.long))
    with pytest.raises(ValueError, match="block size is only"):
        model.max_seq_length = 17
    with pytest.raises(ValueError, match="not a supported config name"):
        Config.from_name("nope")


def test_config_table():
    for name, (L, d, H, G, I, V, hs) in {
        "pythia-70m": (6, 512, 8, 8, 2048, 50304, 64),
        "Llama-2-7b-hf": (32, 4096, 32, 32, 11008, 32000, 128),
        "Llama-2-13b-hf": (40, 5120, 40, 40, 13824, 32000, 128),
        "Llama-2-70b-hf": (80, 8192, 64, 8, 28672, 32000, 128),
        "Mistral-7B-v0.1": (32, 4096, 32, 8, 14336, 32000, 128),
    }.items():
        c = Config.from_name(name)
        assert (c.n_layer, c.n_embd, c.n_head, c.n_query_groups, c.intermediate_size, c.padded_vocab_size, c.head_size) == (L, d, H, G, I, V, hs)
    assert Config.from_name("pythia-70m").rope_n_elem == 16


def test_quantized_module_has_no_cpu_path():
    from lit_gpt_b200 import QLinear, QuantPrecision
    from lit_gpt_b200._cabi import B200Error

    lin = QLinear(64, 32, bias=False, mode="nf4")
    with pytest.raises(B200Error, match="no CPU"):
        lin(torch.zeros(1, 64))
    with pytest.raises(B200Error):
        lin.quantize_()
    model = GPT(Config(block_size=16, vocab_size=16, n_layer=1, n_head=2, n_embd=8))
    QuantPrecision("bnb.nf4", torch.bfloat16).convert_module(model)
    assert all(isinstance(m, QLinear) for n, m in model.named_modules() if n.endswith(("attn.attn", "proj", "lm_head", "fc")))
    sd_keys = set(model.state_dict())
    assert "lm_head.weight" in sd_keys and "transformer.h.0.attn.attn.weight" in sd_keys  # plain weight/bias keys
