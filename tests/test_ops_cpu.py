"""torch.library registration of the C-ABI kernels (lit_gpt_b200/ops.py): schemas, mutation annotations and fake (meta)
kernels -- what makes a traced GPT.forward see single b200gpt::* nodes instead of graph breaks (reference
tests/test_model.py:564-575).  No GPU needed: fake kernels run on meta tensors."""
import torch

import lit_gpt_b200.ops  # noqa: F401  (registers the ops)


def test_ops_registered_with_schemas():
    ns = torch.ops.b200gpt
    for name in ("qlinear", "qlinear_tiled", "rmsnorm", "attn_decode", "rope_kv_append", "attn_prefill", "sample"):
        assert hasattr(ns, name), name
    s = str(ns.attn_decode.default._schema)
    assert "Tensor(a" in s and "k_cache" in s and "v_cache" in s, s      # caches are declared mutable
    assert "Tensor(a" not in str(ns.qlinear.default._schema)              # functional
    assert "rng_state" in str(ns.sample.default._schema) and "Tensor(a" in str(ns.sample.default._schema)


def test_fake_kernels_shapes_on_meta():
    bf = dict(dtype=torch.bfloat16, device="meta")
    ns = torch.ops.b200gpt
    x = torch.empty(2, 5, 4096, **bf)
    packed = torch.empty(11008 * 4096 // 2, 1, dtype=torch.uint8, device="meta")
    absmax = torch.empty(11008 * 4096 // 64, dtype=torch.float32, device="meta")
    y = ns.qlinear(x, packed, absmax, 11008, 4096, "nf4", 64, None)
    assert y.shape == (2, 5, 11008) and y.dtype == torch.bfloat16
    assert ns.qlinear_tiled(x[:, :1], packed, 11008, 4096, "nf4", 64, None).shape == (2, 1, 11008)
    assert ns.rmsnorm(x, torch.empty(4096, **bf), 1e-5).shape == x.shape
    H, G, hs, Tmax = 32, 8, 128, 64
    kc = torch.empty(2, G, Tmax, hs, **bf)
    vc = torch.empty(2, G, Tmax, hs, **bf)
    cos = torch.empty(Tmax, hs, **bf)
    pos1 = torch.empty(1, dtype=torch.int64, device="meta")
    qkv = torch.empty(2, 1, (H + 2 * G) * hs, **bf)
    assert ns.attn_decode(qkv, cos, cos, pos1, kc, vc, H, G, hs, hs, False).shape == (2, H * hs)
    qkvT = torch.empty(2, 5, (H + 2 * G) * hs, **bf)
    posT = torch.empty(5, dtype=torch.int64, device="meta")
    q = ns.rope_kv_append(qkvT, cos, cos, posT, kc, vc, H, G, hs, hs, False)
    assert q.shape == (2, H, 5, hs)
    assert ns.attn_prefill(q, kc, vc, posT, hs ** -0.5).shape == (2, 5, H * hs)
    tok = ns.sample(torch.empty(1, 1, 32000, **bf), 0.8, 200, torch.empty(2, dtype=torch.int64, device="meta"))
    assert tok.shape == (1,) and tok.dtype == torch.int64
