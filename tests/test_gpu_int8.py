"""bnb.int8 (LLM.int8()) CUDA path against oracle/quant_ref.py:310-345.

Integer work (CB, SCB, CA, the int32 products, the rounding chain without outliers) is compared bit-exactly.  With
outlier columns the fp32 side product is summed in ascending column order on the GPU and by a BLAS matmul in the
oracle, so there the tolerance is one fp16 ulp of the result (2^-10 relative + 2^-14 absolute), written below.
"""
import numpy as np
import pytest
import torch

from oracle import quant_ref as Q

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def K():
    from lit_gpt_b200 import kernels

    return kernels


def _rand(shape, seed, std=1.0, dtype=torch.float16):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(shape, generator=g) * std).to(dtype)


@pytest.mark.parametrize("shape,dtype", [((256, 512), torch.float16), ((7, 33), torch.float32), ((64, 11008), torch.bfloat16),
                                         ((1, 1), torch.float16), ((4096, 1024), torch.float16)])
def test_quantize_int8_bit_exact(K, shape, dtype):
    w = _rand(shape, sum(shape), 0.02, dtype)
    if shape[0] > 2:
        w[1] = 0  # all-zero row: SCB 0, CB 0 (np.nan_to_num)
    cb, scb = K.quantize_int8(w.cuda())
    cb_ref, scb_ref = Q.quantize_int8_rowwise(w.float().numpy())
    assert cb.dtype == torch.int8 and scb.dtype == torch.float32
    assert torch.equal(scb.cpu(), torch.from_numpy(scb_ref))
    assert torch.equal(cb.cpu(), torch.from_numpy(cb_ref))
    deq = K.dequantize_int8(cb, scb, torch.float32)
    assert torch.equal(deq.cpu(), torch.from_numpy(Q.dequantize_int8_rowwise(cb_ref, scb_ref)))


def _check(K, M, N, Kd, seed, bias=False, outliers=(), threshold=6.0):
    w = _rand((N, Kd), seed, 0.05)
    x = _rand((M, Kd), seed + 1, 1.0)
    x.clamp_(-5.5, 5.5)
    for (m, k, v) in outliers:
        x[m % M, k % Kd] = v
    b = _rand((N,), seed + 2, 0.1) if bias else None
    cb, scb = K.quantize_int8(w.cuda())
    y = K.linear_int8(x.cuda(), cb, scb, None if b is None else b.cuda(), threshold)
    assert y.shape == (M, N) and y.dtype == torch.float16
    ref = Q.linear_int8(x.float().numpy(), cb.cpu().numpy(), scb.cpu().numpy(), None if b is None else b.float().numpy(), threshold)
    got = y.float().cpu().numpy()
    has_out = threshold > 0 and bool((np.abs(x.float().numpy()) >= threshold).any())
    if has_out:
        assert np.all(np.abs(got - ref) <= np.abs(ref) * 2.0 ** -10 + 2.0 ** -14), np.abs(got - ref).max()
    else:
        assert np.array_equal(got, ref), np.abs(got - ref).max()
    return has_out


@pytest.mark.parametrize("M", [1, 5, 8, 9, 16])
@pytest.mark.parametrize("N,Kd", [(16, 64), (40, 512), (4096, 4096), (1000, 11008)])
def test_linear_int8_decode_exact(K, M, N, Kd):
    assert not _check(K, M, N, Kd, 100 + M + N)


@pytest.mark.parametrize("M", [17, 64, 100, 333])
@pytest.mark.parametrize("N,Kd", [(128, 256), (200, 512), (1024, 4096)])
def test_linear_int8_prefill_exact(K, M, N, Kd):
    assert not _check(K, M, N, Kd, 200 + M + N, bias=(M % 2 == 0))


@pytest.mark.parametrize("M", [1, 4, 16, 70])
def test_linear_int8_outliers_and_bias(K, M):
    out = [(0, 3, 7.5), (1, 3, -6.0), (2, 100, 9.0), (3, 511, -20.0), (0, 64, 6.0)]
    assert _check(K, M, 256, 512, 300 + M, bias=True, outliers=out)
    # threshold 0 disables the side path (bitsandbytes threshold=0.0): large values stay in the int8 product
    assert not _check(K, M, 256, 512, 310 + M, bias=False, outliers=out, threshold=0.0)
    # every column an outlier: the int8 part is identically zero
    assert _check(K, M, 64, 64, 320 + M, outliers=[(0, k, 6.5) for k in range(64)])


def test_linear_int8_zero_and_errors(K):
    from lit_gpt_b200._cabi import B200Error

    cb, scb = K.quantize_int8(_rand((32, 128), 1, 0.05).cuda())
    y = K.linear_int8(torch.zeros(3, 128, dtype=torch.float16, device="cuda"), cb, scb)
    assert torch.count_nonzero(y) == 0 and not torch.isnan(y).any()
    assert K.linear_int8(torch.zeros(0, 128, dtype=torch.float16, device="cuda"), cb, scb).shape == (0, 32)
    with pytest.raises(B200Error):
        K.linear_int8(torch.zeros(3, 128, dtype=torch.bfloat16, device="cuda"), cb, scb)   # fp16 only
    with pytest.raises(B200Error):
        K.linear_int8(torch.zeros(3, 64, dtype=torch.float16, device="cuda"), cb, scb)     # feature mismatch
    cb2, scb2 = K.quantize_int8(_rand((32, 96), 2, 0.05).cuda())
    with pytest.raises(NotImplementedError):
        K.linear_int8(torch.zeros(3, 96, dtype=torch.float16, device="cuda"), cb2, scb2)    # K % 64 != 0


def test_qlinear_int8_module(K):
    """--quantize bnb.int8: QuantPrecision installs QLinear(mode='int8'); weight holds CB, SCB the row statistics."""
    from lit_gpt_b200.quant import QLinear, QuantPrecision

    lin = torch.nn.Linear(512, 256, bias=True)
    with torch.no_grad():
        lin.weight.copy_(_rand((256, 512), 5, 0.05).float())
        lin.bias.copy_(_rand((256,), 6, 0.1).float())
    mod = torch.nn.Sequential(lin).half()
    plugin = QuantPrecision("bnb.int8", torch.float16)
    mod = plugin.convert_module(mod).cuda()
    q = mod[0]
    assert isinstance(q, QLinear) and q.is_quantized and q.weight.dtype == torch.int8 and q.SCB.dtype == torch.float32
    assert tuple(q.quant_state[1]) == (256, 512)   # lit_gpt/utils.py:85-87 reads the original shape at index 1
    cb_ref, scb_ref = Q.quantize_int8_rowwise(lin.weight.detach().half().float().numpy())
    assert torch.equal(q.weight.data.cpu(), torch.from_numpy(cb_ref))
    for M in (1, 40):
        x = _rand((M, 512), 7 + M)
        x[0, 9] = 8.0
        ref = Q.linear_int8(x.float().numpy(), cb_ref, scb_ref, lin.bias.detach().half().float().numpy())
        got = q(x.cuda()).float().cpu().numpy()
        assert np.all(np.abs(got - ref) <= np.abs(ref) * 2.0 ** -10 + 2.0 ** -14)
        # bf16 activations are cast to fp16 for the product and back (bitsandbytes does the same, with a warning)
        yb = q(x.cuda().to(torch.bfloat16))
        assert yb.dtype == torch.bfloat16
    # 3-D input keeps its leading dimensions
    assert q(_rand((2, 3, 512), 11).cuda()).shape == (2, 3, 256)
    # in_features not a multiple of 64: dequantised fp16 product (nothing is silently skipped)
    q2 = QLinear.from_linear(torch.nn.Linear(96, 32, bias=False).half().cuda(), "int8", torch.float16)
    x2 = _rand((2, 96), 12).cuda()
    assert torch.allclose(q2(x2), torch.nn.functional.linear(x2, q2.dequantize(torch.float16)), atol=1e-3, rtol=1e-3)


def test_gpt_int8_generate(K):
    """A small Llama-shaped GPT under bnb.int8 / 16-true: logits stay close to the fp16 model, generate() runs."""
    from lit_gpt_b200 import GPT, Config, QuantPrecision
    from lit_gpt_b200.fabric import Fabric
    from lit_gpt_b200.generate.base import generate

    cfg = dict(block_size=64, vocab_size=320, padded_vocab_size=320, n_layer=2, n_head=4, n_embd=128, n_query_groups=2,
               rotary_percentage=1.0, parallel_residual=False, bias=False, _norm_class="RMSNorm", _mlp_class="LLaMAMLP",
               intermediate_size=256)
    torch.manual_seed(0)
    ref = GPT(Config(**cfg)).half().cuda().eval()
    ref.max_seq_length = 32
    ref.set_kv_cache(batch_size=1, device="cuda", dtype=torch.float16)
    sd = {k: v.clone() for k, v in ref.state_dict().items()}

    fabric = Fabric(devices=1, plugins=QuantPrecision("bnb.int8", torch.float16))
    with fabric.init_module(empty_init=True):
        model = GPT(Config(**cfg))
    with fabric.init_tensor():
        model.max_seq_length = 32
        model.set_kv_cache(batch_size=1)
    model.eval()
    model = fabric.setup_module(model)
    model.load_state_dict(sd, strict=True)
    from lit_gpt_b200.quant import QLinear

    assert all(isinstance(m, QLinear) and m.quant_type == "int8" for m in (model.lm_head, model.transformer.h[0].attn.attn))
    idx = torch.randint(0, 320, (1, 8), device="cuda")
    pos = torch.arange(8, device="cuda")
    with torch.no_grad():
        a = ref(idx, pos).float()
        b = model(idx, pos).float()
    assert b.shape == a.shape and torch.isfinite(b).all()
    assert (a - b).abs().max() <= 0.05 * a.abs().max() + 0.05   # 8-bit weights and activations: ~1 % relative noise
    with torch.no_grad():
        out = generate(model, idx[0], 16, temperature=0.0)
    assert out.shape == (16,) and torch.equal(out[:8], idx[0])
