"""Pre-quantized checkpoint round trip (lit_gpt_b200/checkpoint.py): what is written is what sits in HBM, and a
model loaded from it -- without running the quantizer -- is bit-identical to the one that was quantized from 16-bit weights."""
import pytest
import torch

from oracle import model_ref as M

pytestmark = pytest.mark.gpu

KW = dict(name="ckpt", block_size=64, vocab_size=512, padded_vocab_size=512, n_layer=2, n_head=4, n_query_groups=2, n_embd=256,
          rotary_percentage=1.0, parallel_residual=False, bias=False, _norm_class="RMSNorm", _mlp_class="LLaMAMLP", intermediate_size=512)


@pytest.mark.parametrize("mode", ["nf4", "nf4-dq", "fp4", "int4", "int8"])
@torch.inference_mode()
def test_save_load_quantized_bit_identical(mode, tmp_path, monkeypatch):
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from test_gpu_model import build_cuda_model

    from lit_gpt_b200 import GPT, Config, QLinear, QuantPrecision
    from lit_gpt_b200 import kernels as K
    from lit_gpt_b200.checkpoint import load_quantized, save_quantized
    from lit_gpt_b200.fabric import Fabric

    dtype = torch.float16 if mode == "int8" else torch.bfloat16
    p = M.init_weights(M.RefConfig.from_any(Config(**KW)), seed=11, dtype=torch.bfloat16)
    a = build_cuda_model(KW, p, mode, max_seq=32)
    path = tmp_path / "model.q.pth"
    save_quantized(a, path)

    fabric = Fabric(devices=1, plugins=QuantPrecision(mode, torch.bfloat16))
    with fabric.init_module(empty_init=True):
        b = GPT(Config(**KW))
    with fabric.init_tensor():
        b.max_seq_length = 32
        b.set_kv_cache(batch_size=1)
    b.eval()
    # loading must not quantize anything
    def boom(*a_, **k_):
        raise AssertionError("the quantizer ran while loading a pre-quantized checkpoint")
    for fn in ("quantize_4bit", "quantize_int4", "quantize_int8", "double_quant_absmax"):
        monkeypatch.setattr(K, fn, boom)
    load_quantized(b, path, "cuda")

    for (na, ma), (nb, mb) in zip(a.named_modules(), b.named_modules()):
        if isinstance(ma, QLinear):
            assert na == nb and isinstance(mb, QLinear) and mb.is_quantized
            assert torch.equal(ma.weight.data, mb.weight.data), na
            for attr in ("absmax", "scales", "qzeros", "sz", "SCB", "tiled"):
                ta, tb = getattr(ma, attr), getattr(mb, attr)
                assert (ta is None) == (tb is None), (na, attr)
                if ta is not None:
                    assert torch.equal(ta, tb), (na, attr)
            assert tuple(ma.quant_state[1]) == tuple(mb.quant_state[1]) and ma.quant_state[5] == mb.quant_state[5]
    assert torch.equal(a.transformer.wte.weight, b.transformer.wte.weight)
    assert torch.equal(a.transformer.ln_f.weight, b.transformer.ln_f.weight)

    prompt = torch.tensor([[3, 1, 4, 1, 5, 9]], device="cuda")
    outs = []
    for m in (a, b):
        for blk in m.transformer.h:
            blk.attn.kv_cache.reset_parameters()
        y = m(prompt, torch.arange(6, device="cuda"))
        d = m(torch.tensor([[7]], device="cuda"), torch.tensor([6], device="cuda"))
        outs.append((y.clone(), d.clone()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])


def test_load_quantized_rejects_mismatch(tmp_path):
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from test_gpu_model import build_cuda_model

    from lit_gpt_b200 import Config
    from lit_gpt_b200._cabi import B200Error
    from lit_gpt_b200.checkpoint import load_quantized, save_quantized

    p = M.init_weights(M.RefConfig.from_any(Config(**KW)), seed=11, dtype=torch.bfloat16)
    a = build_cuda_model(KW, p, "nf4", max_seq=16)
    save_quantized(a, tmp_path / "a.pth")
    b = build_cuda_model(KW, p, "int4", max_seq=16)
    with pytest.raises(B200Error, match="mode"):
        load_quantized(b, tmp_path / "a.pth")
    torch.save({"format": "something else"}, tmp_path / "bad.pth")
    with pytest.raises(B200Error, match="not a"):
        load_quantized(b, tmp_path / "bad.pth")
