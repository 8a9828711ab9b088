"""Pins oracle/model_ref.py against golden vectors produced by the unmodified reference (tests/golden)."""
import ast
import hashlib

import numpy as np
import pytest
import torch

from oracle import model_ref as M
from oracle.ref_loader import load_reference, reference_available

CASES = ["llama2_toy", "llama2_70b_mqa_toy", "mistral_toy", "llama_gqa_hs64", "llama_mha_hs128", "pythia_toy"]


def _sha(p):
    h = hashlib.sha256()
    for k in sorted(p):
        h.update(k.encode())
        h.update(p[k].detach().contiguous().numpy().tobytes())
    return h.hexdigest()


def load_case(golden, name):
    g = golden(name)
    cfgd = ast.literal_eval(str(g["config"]))
    rc = M.RefConfig.from_any(cfgd)
    p = M.init_weights(rc, seed=1234, bias=cfgd["bias"])
    if cfgd["_norm_class"] == "LayerNorm":
        gen = torch.Generator().manual_seed(7)
        for k in p:
            if "norm" in k or "ln_f" in k:
                p[k] = p[k] + 0.1 * torch.randn(p[k].shape, generator=gen)
    assert _sha(p) == str(g["sha"]), "seeded weight generation drifted from the golden fixture"
    return g, cfgd, rc, p


@pytest.mark.parametrize("name", CASES)
@torch.inference_mode()
def test_forward_matches_reference_golden(golden, name):
    g, _, rc, p = load_case(golden, name)
    x = torch.from_numpy(g["fwd_input"])
    cos, sin = M.build_rope_cache(rc.block_size, rc.rope_n_elem, rc.rope_base, rc.rope_condense_ratio)
    y = M.gpt_forward(rc, p, x, cos=cos, sin=sin)
    torch.testing.assert_close(y, torch.from_numpy(g["fwd_logits"]))


@pytest.mark.parametrize("name", CASES)
@torch.inference_mode()
def test_generate_matches_reference_golden(golden, name):
    g, _, rc, p = load_case(golden, name)
    prompt = torch.from_numpy(g["gen_prompt"])
    n_total = len(g["gen_tokens"])
    st = M.RefState(rc, n_total, 1, torch.float32)
    toks, logits = M.generate(rc, p, st, prompt, n_total, temperature=0.0, return_logits=True)
    assert toks.tolist() == g["gen_tokens"].tolist()
    torch.testing.assert_close(logits, torch.from_numpy(g["gen_logits"]))


@pytest.mark.parametrize("name", CASES)
@torch.inference_mode()
def test_batch2_kv_decode_matches_reference_golden(golden, name):
    g, _, rc, p = load_case(golden, name)
    x = torch.from_numpy(g["fwd_input"])
    xb = torch.stack([x[0], x[0].flip(0)])
    st = M.RefState(rc, 25, 2, torch.float32)
    y = M.gpt_forward(rc, p, xb, torch.arange(5), st)
    torch.testing.assert_close(y, torch.from_numpy(g["b2_prefill_logits"]))
    y = M.gpt_forward(rc, p, torch.tensor([[3], [7]]), torch.tensor([5]), st)
    torch.testing.assert_close(y, torch.from_numpy(g["b2_decode_logits"]))


@torch.inference_mode()
def test_units(golden):
    u = golden("units")
    x, w = torch.from_numpy(u["rms_x"]), torch.from_numpy(u["rms_w"])
    torch.testing.assert_close(M.rmsnorm(x, w, 1e-5), torch.from_numpy(u["rms_y"]))
    assert torch.equal(M.rmsnorm(x.bfloat16(), w, 1e-5).float(), torch.from_numpy(u["rms_y_bf16"]))
    rows = torch.from_numpy(u["rope_rows"])
    cos, sin = M.build_rope_cache(4096, 128, pos_dtype=torch.float32)
    torch.testing.assert_close(cos[rows], torch.from_numpy(u["rope_cos_f32"]))
    torch.testing.assert_close(sin[rows], torch.from_numpy(u["rope_sin_f32"]))
    cosb, sinb = M.build_rope_cache(4096, 128, pos_dtype=torch.bfloat16)
    torch.testing.assert_close(cosb[rows], torch.from_numpy(u["rope_cos_bf16pos"]))
    assert len(torch.unique(cosb, dim=0)) == int(u["rope_bf16pos_distinct"]) == 769
    q = torch.from_numpy(u["rope_x"])
    torch.testing.assert_close(M.apply_rope(q, cos[:7], sin[:7]), torch.from_numpy(u["rope_y"]))
    assert torch.equal(M.apply_rope(q.bfloat16(), cos[:7], sin[:7]).float(), torch.from_numpy(u["rope_y_bf16"]))
    assert torch.equal(M.apply_rope(q.bfloat16(), cos[:7].bfloat16(), sin[:7].bfloat16()).float(),
                       torch.from_numpy(u["rope_y_bf16_tab16"]))
    w8 = torch.arange(64, dtype=torch.float32).reshape(8, 8)
    assert torch.equal(M.tensor_parallel_shard(w8, 4, 2, "colwise"), torch.from_numpy(u["tp_col_w"]))
    assert torch.equal(M.tensor_parallel_shard(w8, 4, 2, "rowwise"), torch.from_numpy(u["tp_row_w"]))
    assert M.sample(torch.from_numpy(u["sample_logits"]), temperature=0.0).tolist() == u["sample_t0"].tolist() == [0]


@pytest.mark.skipif(not reference_available(), reason="/root/reference not present (GPU box)")
@torch.inference_mode()
def test_against_live_reference_bf16(golden):
    """bf16 rounding points of the restatement equal the reference's, op by op (bit-exact on CPU)."""
    R = load_reference()
    _, cfgd, rc, p = load_case(golden, "llama_gqa_hs64")
    cfg = R.Config(**{k: v for k, v in cfgd.items() if k not in ("head_size", "rope_n_elem")},
                   rotary_percentage=cfgd["rope_n_elem"] / cfgd["head_size"])
    pb = {k: v.bfloat16() for k, v in p.items()}
    torch.set_default_dtype(torch.bfloat16)
    model = R.GPT(cfg)
    torch.set_default_dtype(torch.float32)
    model.load_state_dict(pb)
    model.max_seq_length = 30
    model.set_kv_cache(1, dtype=torch.bfloat16)
    st = M.RefState(rc, 30, 1, torch.bfloat16, cos=model.cos, sin=model.sin)
    prompt = torch.tensor([3, 1, 4, 1, 5, 9, 2, 6])
    y_ref = model(prompt.view(1, -1), torch.arange(8))
    y = M.gpt_forward(rc, pb, prompt.view(1, -1), torch.arange(8), st)
    assert torch.equal(y, y_ref)
    y_ref = model(torch.tensor([[7]]), torch.tensor([8]))
    y = M.gpt_forward(rc, pb, torch.tensor([[7]]), torch.tensor([8]), st)
    assert torch.equal(y, y_ref)
