"""Generate golden vectors by running the UNMODIFIED reference (/root/reference) in this container.

    python tests/golden/make_golden.py            # rewrites tests/golden/*.npz

The reference cannot travel to the GPU box, so its outputs are committed as small fixtures.  Weights are
produced by ``oracle.model_ref.init_weights(seed)`` (a seeded torch CPU generator) and loaded into the
reference's own ``GPT`` with ``load_state_dict``; the fixture stores a sha256 of the weights so RNG drift is
detected, plus inputs and the reference's logits / tokens.  Reference entry points exercised:
lit_gpt/model.py:499-519 (GPT.forward), 534-554 (set_kv_cache), generate/base.py:50-93 (generate),
lit_gpt/rmsnorm.py:19-25, lit_gpt/model.py:746-773 (rope), generate/tp.py:28-45 (tensor_parallel_linear).
"""
from __future__ import annotations

import hashlib
import os
import sys
from unittest.mock import Mock

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.abspath(os.path.join(HERE, "..", "..")))

from oracle import model_ref as M  # noqa: E402
from oracle.ref_loader import load_reference  # noqa: E402

CASES = {
    # name -> (Config.from_name kwargs); first three mirror tests/test_model.py:201-261,403-447 of the reference
    "llama2_toy": dict(name="Llama-2-7b-hf", padded_vocab_size=1024, n_layer=2, n_head=8, n_embd=32, intermediate_size=86),
    "llama2_70b_mqa_toy": dict(name="Llama-2-70b-chat-hf", n_query_groups=1, padded_vocab_size=1024, n_layer=2, n_head=8, n_embd=32, intermediate_size=86),
    "mistral_toy": dict(name="Mistral-7B-Instruct-v0.1", padded_vocab_size=1024, n_layer=2, n_embd=32, n_head=8, n_query_groups=2, intermediate_size=86),
    # head_size 64 / 128 shaped cases used for kv-cache decode goldens
    "llama_gqa_hs64": dict(name="Llama-2-7b-hf", padded_vocab_size=512, n_layer=2, n_head=4, n_query_groups=2, n_embd=256, intermediate_size=352, block_size=64),
    "llama_mha_hs128": dict(name="Llama-2-7b-hf", padded_vocab_size=512, n_layer=2, n_head=2, n_embd=256, intermediate_size=384, block_size=64),
    # BASELINE config 1 family: LayerNorm + GELU + bias + parallel residual + partial rotary
    "pythia_toy": dict(name="pythia-14m", n_layer=2, block_size=64, padded_vocab_size=1024),
}
FIXED_INPUT = [[9856, 23, 491, 1536, 304]]  # reference tests/test_model.py:256


def weights_sha(p) -> str:
    h = hashlib.sha256()
    for k in sorted(p):
        h.update(k.encode())
        h.update(p[k].detach().contiguous().numpy().tobytes())
    return h.hexdigest()


def config_dict(cfg) -> dict:
    keys = ["n_layer", "n_head", "n_embd", "n_query_groups", "head_size", "rope_n_elem", "intermediate_size",
            "padded_vocab_size", "block_size", "norm_eps", "rope_base", "rope_condense_ratio", "parallel_residual",
            "shared_attention_norm", "_norm_class", "_mlp_class", "gelu_approximate", "bias", "vocab_size"]
    return {k: getattr(cfg, k) for k in keys}


@torch.inference_mode()
def main() -> None:
    torch.set_default_dtype(torch.float32)
    R = load_reference()
    for name, kwargs in CASES.items():
        kwargs = dict(kwargs)
        cfg = R.Config.from_name(kwargs.pop("name"), **kwargs)
        rc = M.RefConfig.from_any(cfg)
        p = M.init_weights(rc, seed=1234, bias=cfg.bias)
        if cfg._norm_class == "LayerNorm":  # make the norm affine non-trivial
            g = torch.Generator().manual_seed(7)
            for k in p:
                if "norm" in k or "ln_f" in k:
                    p[k] = p[k] + 0.1 * torch.randn(p[k].shape, generator=g)
        model = R.GPT(cfg)
        model.load_state_dict(p, strict=True)
        model.eval()
        out = {"sha": np.array(weights_sha(p)), "config": np.array(repr(config_dict(cfg)))}
        V = cfg.padded_vocab_size
        x = torch.tensor(FIXED_INPUT) % V
        out["fwd_input"] = x.numpy()
        out["fwd_logits"] = model(x).numpy()
        # kv-cache path: prefill 5 tokens then greedy decode, through the reference's own generate()
        T, new = 5, 20
        model.max_seq_length = T + new
        model.set_kv_cache(batch_size=1)
        logits_log = []
        orig_forward = model.forward

        def spy(idx, input_pos=None):
            y = orig_forward(idx, input_pos)
            logits_log.append(y[0, -1].clone())
            return y

        model.forward = spy
        prompt = x[0]
        toks = R.generate(model, prompt, T + new, temperature=0.0)
        model.forward = orig_forward
        out["gen_prompt"] = prompt.numpy()
        out["gen_tokens"] = toks.numpy()
        out["gen_logits"] = torch.stack(logits_log).numpy()
        # batch-2 decode step at the model level (generate() is batch-1 only)
        model.set_kv_cache(batch_size=2)
        xb = torch.stack([x[0], x[0].flip(0)])
        out["b2_prefill_logits"] = model(xb, torch.arange(T)).numpy()
        out["b2_decode_logits"] = model(torch.tensor([[3], [7]]), torch.tensor([T])).numpy()
        np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **out)
        print(name, "logits", out["fwd_logits"].shape, "tokens", out["gen_tokens"][-5:])

    # unit vectors: RMSNorm, rope tables (fp32 and the bf16-default-dtype aliasing of SURVEY F5), apply_rope
    g = torch.Generator().manual_seed(99)
    u = {}
    xx = torch.randn(3, 7, 96, generator=g)
    w = torch.randn(96, generator=g)
    norm = R.RMSNorm(96, eps=1e-5)
    norm.weight.data = w
    u["rms_x"], u["rms_w"], u["rms_y"] = xx.numpy(), w.numpy(), norm(xx).numpy()
    u["rms_y_bf16"] = norm(xx.bfloat16()).float().numpy()
    cos, sin = R.model.build_rope_cache(4096, 128)
    u["rope_rows"] = np.array([0, 1, 255, 256, 257, 2047, 4095])
    u["rope_cos_f32"], u["rope_sin_f32"] = cos[u["rope_rows"]].numpy(), sin[u["rope_rows"]].numpy()
    torch.set_default_dtype(torch.bfloat16)
    cosb, sinb = R.model.build_rope_cache(4096, 128)
    torch.set_default_dtype(torch.float32)
    assert cosb.dtype == torch.float32
    u["rope_cos_bf16pos"], u["rope_sin_bf16pos"] = cosb[u["rope_rows"]].numpy(), sinb[u["rope_rows"]].numpy()
    u["rope_bf16pos_distinct"] = np.array(len(torch.unique(cosb, dim=0)))
    q = torch.randn(2, 4, 7, 128, generator=g)
    u["rope_x"] = q.numpy()
    u["rope_y"] = R.model.apply_rope(q, cos[:7], sin[:7]).numpy()
    u["rope_y_bf16"] = R.model.apply_rope(q.bfloat16(), cos[:7], sin[:7]).float().numpy()
    u["rope_y_bf16_tab16"] = R.model.apply_rope(q.bfloat16(), cos[:7].bfloat16(), sin[:7].bfloat16()).float().numpy()
    # tensor_parallel_linear (reference tests/test_generate_tp.py:14-40)
    fabric = Mock()
    fabric.world_size, fabric.global_rank = 4, 2
    lin = torch.nn.Linear(8, 8, bias=True)
    lin.weight.data = torch.arange(64, dtype=torch.float32).reshape(8, 8)
    lin.bias.data = torch.arange(8, dtype=torch.float32)
    R.tensor_parallel_linear(fabric, lin, "colwise")
    u["tp_col_w"], u["tp_col_b"] = lin.weight.detach().numpy(), lin.bias.detach().numpy()
    lin = torch.nn.Linear(8, 8, bias=False)
    lin.weight.data = torch.arange(64, dtype=torch.float32).reshape(8, 8)
    R.tensor_parallel_linear(fabric, lin, "rowwise")
    u["tp_row_w"] = lin.weight.detach().numpy()
    # sample (reference tests/test_generate.py:94-107)
    logits = torch.tensor([[[24, 4, 98, 77, 47], [65, 70, 32, 67, 24], [92, 32, 88, 36, 62]],
                           [[85, 79, 57, 68, 50], [89, 46, 72, 45, 32], [68, 96, 68, 24, 36]]])
    u["sample_logits"] = logits.numpy()
    u["sample_t0"] = R.sample(logits, temperature=0.0).numpy()
    np.savez_compressed(os.path.join(HERE, "units.npz"), **u)
    print("units ok; bf16-position rope table has", int(u["rope_bf16pos_distinct"]), "distinct rows of 4096")


if __name__ == "__main__":
    main()
