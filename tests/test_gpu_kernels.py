"""Parity of each CUDA kernel (through the C ABI) against the oracle, on seeded inputs.  Run with -m gpu on a B200."""
import math

import numpy as np
import pytest
import torch

from oracle import model_ref as M
from oracle import quant_ref as Q
from oracle import quant_ref_c as QC

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def K():
    from lit_gpt_b200 import kernels

    return kernels


def _randw(shape, seed, dtype=torch.bfloat16, std=0.02):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(shape, generator=g) * std).to(dtype)


# ---------------------------------------------------------------------------------------------------------
# quantise / dequantise: bit-exact
# ---------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("qt", ["nf4", "fp4"])
@pytest.mark.parametrize("shape,dtype", [((256, 512), torch.bfloat16), ((7, 33), torch.float32), ((1, 64), torch.float16),
                                         ((3, 1), torch.bfloat16), ((4096, 1024), torch.bfloat16), ((64, 11008), torch.bfloat16)])
def test_quantize_4bit_bit_exact(K, qt, shape, dtype):
    w = _randw(shape, seed=sum(shape), dtype=dtype)
    w[0, 0] = 0.0
    packed, absmax = K.quantize_4bit(w.cuda(), qt)
    p_ref, a_ref = QC.quantize_4bit(w.float().numpy(), qt)
    assert packed.shape == p_ref.shape and packed.dtype == torch.uint8
    assert np.array_equal(packed.cpu().numpy(), p_ref)
    assert np.array_equal(absmax.cpu().numpy(), a_ref)
    d32 = K.dequantize_4bit(packed, absmax, shape, torch.float32, qt)
    d_ref = QC.dequantize_4bit(p_ref, a_ref, shape, qt)
    assert np.array_equal(d32.cpu().numpy(), d_ref)
    d16 = K.dequantize_4bit(packed, absmax, shape, torch.bfloat16, qt)
    assert torch.equal(d16.cpu(), torch.from_numpy(d_ref).to(torch.bfloat16))
    # idempotence through the device: quantise(dequantise) reproduces absmax and the dequantised values
    p2, a2 = K.quantize_4bit(d32, qt)
    assert torch.equal(a2, absmax) and torch.equal(K.dequantize_4bit(p2, a2, shape, torch.float32, qt), d32)


def test_quantize_4bit_edge_blocks(K):
    w = torch.zeros(3, 64)
    w[1, 5], w[2, :] = 2.5, -1e-30  # all-zero block, single spike, denormal-scale block
    packed, absmax = K.quantize_4bit(w.cuda(), "nf4")
    p_ref, a_ref = Q.quantize_4bit(w.numpy(), "nf4")
    assert np.array_equal(packed.cpu().numpy(), p_ref) and np.array_equal(absmax.cpu().numpy(), a_ref)
    assert torch.all(K.dequantize_4bit(packed, absmax, (3, 64), torch.float32)[0] == 0)


@pytest.mark.parametrize("shape", [(64, 256), (5, 200), (16, 4096), (3, 8), (32, 11008)])
def test_int4_bit_exact(K, shape):
    w = _randw(shape, seed=shape[1])
    qw, scales, qz, sz = K.quantize_int4(w.cuda(), 128)
    qw_ref, s_ref, z_ref = Q.quantize_int4_g128(w.float().numpy())
    assert np.array_equal(qw.cpu().numpy().view(np.uint32), qw_ref)
    assert np.array_equal(scales.cpu().numpy(), s_ref)
    assert np.array_equal(qz.cpu().numpy().view(np.uint32), z_ref)
    d = K.dequantize_int4(qw, scales, qz, shape[1], 128, torch.float32)
    assert np.array_equal(d.cpu().numpy(), QC.dequantize_int4(qw_ref, s_ref, z_ref, shape[1]))
    # kernel-side {scale, zero} pairs hold the same information
    szn = sz.cpu().numpy().view(np.uint32)
    assert np.array_equal((szn & 0xFFFF).astype(np.uint16).view(np.float16), s_ref)


# ---------------------------------------------------------------------------------------------------------
# fused dequant-GEMV
# ---------------------------------------------------------------------------------------------------------
def _gemv_case(K, N, Kdim, B, qt, seed=0, **kw):
    w = _randw((N, Kdim), seed)
    x = _randw((B, Kdim), seed + 1, std=1.0)
    wd = w.cuda()
    if qt == "int4":
        qw, scales, qz, sz = K.quantize_int4(wd, 128)
        packed, scale, bs = qw, sz, 128
        deq = QC.dequantize_int4(qw.cpu().numpy().view(np.uint32), scales.cpu().numpy(), qz.cpu().numpy().view(np.uint32), Kdim)
    else:
        packed, scale = K.quantize_4bit(wd, qt)
        bs = 64
        deq = QC.dequantize_4bit(packed.cpu().numpy(), scale.cpu().numpy(), (N, Kdim), qt)
    return w, x, packed, scale, bs, torch.from_numpy(deq)


def _check_gemv(y, x, deq, extra=None, post=None):
    """kernel error against the wide oracle must stay within 2x the error of the plain bf16 torch composition."""
    wide = x.double() @ deq.double().T
    comp = torch.nn.functional.linear(x.to(torch.bfloat16).cuda(), deq.to(torch.bfloat16).cuda()).float().cpu()  # reference's own formulation
    if post is not None:
        wide, comp = post(wide.float()).double(), post(comp)
    e_k = (y.float().cpu().double() - wide).abs().max().item()
    e_t = (comp.double() - wide).abs().max().item()
    scale = wide.abs().max().item()
    assert e_k <= 2.0 * e_t + 2 ** -8 * scale, f"kernel err {e_k:.4g} vs torch-bf16 err {e_t:.4g} (|y|max {scale:.3g})"


@pytest.mark.parametrize("qt", ["nf4", "fp4", "int4"])
@pytest.mark.parametrize("N,Kdim,B", [(512, 4096, 1), (4096, 4096, 1), (344, 11008, 1), (1000, 5120, 3), (4096, 1024, 8),
                                      (2048, 4096, 16), (256, 256, 5), (20, 64, 1), (512, 8192, 2)])
def test_qgemv_fast_path(K, qt, N, Kdim, B):
    w, x, packed, scale, bs, deq = _gemv_case(K, N, Kdim, B, qt, seed=N + B)
    y = K.qgemv(x.cuda(), packed, scale, N, Kdim, qt, bs)
    assert y.shape == (B, N) and y.dtype == torch.bfloat16
    _check_gemv(y, x.float(), deq)
    y2 = K.qgemv(x.cuda(), packed, scale, N, Kdim, qt, bs)  # split-K counters are self-cleaning: run twice
    assert torch.equal(y, y2), "not deterministic / workspace not clean between launches"


@pytest.mark.parametrize("ks,occ", [(1, 1), (2, 2), (4, 1), (4, 3)])
def test_qgemv_tuning_variants_agree(K, ks, occ):
    w, x, packed, scale, bs, deq = _gemv_case(K, 1024, 4096, 2, "nf4", seed=5)
    try:
        K.qgemv_tune(ks, occ)
        y = K.qgemv(x.cuda(), packed, scale, 1024, 4096, "nf4", bs)
    finally:
        K.qgemv_tune(0, 0)
    _check_gemv(y, x.float(), deq)


@pytest.mark.parametrize("qt,H,Kdim", [("nf4", 32, 4096), ("int4", 40, 5120)])
@pytest.mark.parametrize("pos", [0, 5, 271, 700])
def test_qgemv_tiled_fused_attention(K, qt, H, Kdim, pos):
    """The decode-attention step fused into the QKV GEMV (b200_qgemv_tiled_args::attn; lit_gpt/model.py:619 + 624-653 in one
    launch) against the two separate launches: identical QKV row and cache append, attention output equal up to the order in
    which the key ranges are merged (fp32 merge, bf16 output: 2^-7 relative)."""
    hs, max_seq = 128, 1024
    N = 3 * H * hs
    w, x, packed, scale, bs, deq = _gemv_case(K, N, Kdim, 1, qt, seed=H + pos)
    tiled = K.tile_4bit(packed, scale, N, Kdim, qt, bs)
    assert K.qgemv_attn_fusable(N, Kdim, qt, H, H, hs, hs, max_seq)
    g = torch.Generator().manual_seed(pos)
    kc0 = torch.randn(1, H, max_seq, hs, generator=g).bfloat16().cuda()
    vc0 = torch.randn(1, H, max_seq, hs, generator=g).bfloat16().cuda()
    cos, sin = M.build_rope_cache(max_seq, hs)
    cos, sin = cos.cuda(), sin.cuda()
    nw = torch.ones(Kdim, dtype=torch.bfloat16, device="cuda")
    p = torch.tensor([pos], device="cuda")
    xs = (x * 40).cuda()   # spread the logits so that the softmax is not flat
    # separate launches
    kc1, vc1 = kc0.clone(), vc0.clone()
    qkv1 = K.qgemv_tiled(xs, tiled, N, Kdim, qt, bs, norm_weight=nw, norm_eps=1e-5)
    y1 = K.attn_decode(qkv1, cos, sin, p, kc1, vc1, H, H, hs, hs)
    # fused
    for rep in range(2):   # twice: the per-head counters must be left clean
        kc2, vc2 = kc0.clone(), vc0.clone()
        y2 = torch.full((1, H * hs), float("nan"), dtype=torch.bfloat16, device="cuda")
        qkv2 = K.qgemv_tiled(xs, tiled, N, Kdim, qt, bs, norm_weight=nw, norm_eps=1e-5,
                             attn=dict(k_cache=kc2, v_cache=vc2, pos=p, cos=cos, sin=sin, out=y2, n_head=H, n_groups=H, head_size=hs, rope_n_elem=hs))
        torch.cuda.synchronize()
        assert torch.equal(qkv1, qkv2)
        assert torch.equal(kc1, kc2) and torch.equal(vc1, vc2)
        assert torch.isfinite(y2.float()).all()
        assert (y1.float() - y2.float()).abs().max() <= 2 ** -7 * y1.float().abs().max(), (y1.float() - y2.float()).abs().max()
    K.check_gemv_errors(torch.device("cuda", 0))


@pytest.mark.parametrize("qt", ["nf4", "int4"])
@pytest.mark.parametrize("N,Kdim,B", [(5120, 13824, 16), (2048, 13824, 8), (1000, 5120, 16), (640, 8192, 16), (5120, 13824, 9), (96, 28672, 16)])
def test_qgemv_tiled_k_windows(K, qt, N, Kdim, B):
    """Batch x K too large for the activation rows to sit in shared memory beside the ring (Llama-2-13B gptq.int4 at batch
    16, BASELINE config 3): the kernel sweeps its tiles once per k-window.  Plain, norm + delta prologue with the residual
    epilogue, and the SwiGLU pair, against the oracle's dequantised weights."""
    w, x, packed, scale, bs, deq = _gemv_case(K, N, Kdim, B, qt, seed=N + B)
    tiled = K.tile_4bit(packed, scale, N, Kdim, qt, bs)
    assert K.qgemv_tiled_supported(N, Kdim, qt, B)
    y = K.qgemv_tiled(x.cuda(), tiled, N, Kdim, qt, bs)
    _check_gemv(y, x.float(), deq)
    assert torch.equal(y, K.qgemv_tiled(x.cuda(), tiled, N, Kdim, qt, bs)), "not deterministic"
    # prologue: (x + delta) -> x_out, RMSNorm; epilogue: residual
    delta, nw, res = _randw((B, Kdim), 3, std=1.0), _randw((Kdim,), 4, std=1.0), _randw((B, N), 5, std=1.0)
    x_out = torch.empty(B, Kdim, dtype=torch.bfloat16, device="cuda")
    y2 = K.qgemv_tiled(x.cuda(), tiled, N, Kdim, qt, bs, delta=delta.cuda(), x_out=x_out, norm_weight=nw.cuda(), norm_eps=1e-5, residual=res.cuda())
    xs = (x + delta)  # bf16 add, as `x = h + x` in the reference (model.py:591)
    assert torch.equal(x_out.cpu(), xs)
    xn = M.rmsnorm(xs, nw, 1e-5)  # oracle rounding points (rmsnorm.py:19-25)
    _check_gemv(y2, xn.float(), deq, post=lambda t: t.to(torch.bfloat16).float() + res.float())


@pytest.mark.parametrize("qt", ["nf4", "int4"])
def test_qgemv_tiled_k_windows_swiglu(K, qt):
    N, Kdim, B = 1376, 8192, 16
    w, x, packed, scale, bs, deq = _gemv_case(K, N, Kdim, B, qt, seed=11)
    w2, _, packed2, scale2, _, deq2 = _gemv_case(K, N, Kdim, B, qt, seed=12)
    t1, t2 = K.tile_4bit(packed, scale, N, Kdim, qt, bs), K.tile_4bit(packed2, scale2, N, Kdim, qt, bs)
    assert K.qgemv_tiled_supported(N, Kdim, qt, B, True)
    y = K.qgemv_tiled(x.cuda(), t1, N, Kdim, qt, bs, tiled2=t2).float().cpu()
    xb = x.to(torch.bfloat16).cuda()
    a = torch.nn.functional.linear(xb, deq.to(torch.bfloat16).cuda())
    b = torch.nn.functional.linear(xb, deq2.to(torch.bfloat16).cuda())
    comp = (torch.nn.functional.silu(a) * b).float().cpu()
    wide = (torch.nn.functional.silu(x.double() @ deq.double().T) * (x.double() @ deq2.double().T)).float()
    e_k, e_t = (y - wide).abs().max().item(), (comp - wide).abs().max().item()
    assert e_k <= 2.0 * e_t + 2 ** -8 * wide.abs().max().item(), (e_k, e_t)


@pytest.mark.parametrize("qt", ["nf4", "fp4", "int4"])
@pytest.mark.parametrize("N,Kdim,B", [(4096, 11008, 16), (5120, 13824, 16), (4096, 4096, 9), (12288, 4096, 16), (1000, 5120, 5), (48, 28672, 2),
                                      (32000, 1024, 16), (4096, 128, 12), (16, 256, 16), (4112, 2048, 7)])
def test_qgemv_batched_tile_per_warp(K, qt, N, Kdim, B):
    """csrc/qgemv_nb.cu: 2..16 PLAIN rows, one 16-row tile per warp, the activation slice streamed with the weights, groups of 16
    tiles cut by a stream-K boundary summed in CTA order through self-validating words in the workspace.  Shapes: many
    contributors per group (4096 x 11008, 5120 x 13824: ~9 CTAs per group), whole groups per CTA (32000 x 1024), a single step per
    tile (K = 128 / 256), partial last groups (1000, 48, 16, 4112 rows), every batch parity.  Checked against the oracle's
    dequantised weights, against the k-split kernel (qgemv_v4.cu, forced with the tuning switch) and for run-to-run equality
    (deterministic order, workspace left clean); RESIDUAL in place (y aliases residual), bias, SwiGLU pair."""
    if qt == "fp4" and (N, Kdim, B) not in ((4096, 4096, 9), (1000, 5120, 5)):
        pytest.skip("fp4 shares the nf4 path: sampled subset")
    w, x, packed, scale, bs, deq = _gemv_case(K, N, Kdim, B, qt, seed=N + Kdim + B)
    tiled = K.tile_4bit(packed, scale, N, Kdim, qt, bs)
    xd = x.cuda()
    n0 = K.launch_count()
    y = K.qgemv_tiled(xd, tiled, N, Kdim, qt, bs)
    assert K.launch_count() == n0 + 1
    _check_gemv(y, x.float(), deq)
    # a different shape in between (its cut groups use the same workspace words), then the same launch again
    w_b, x_b, p_b, s_b, _, deq_b = _gemv_case(K, 4096, 2048, 16, qt, seed=1)
    _check_gemv(K.qgemv_tiled(x_b.cuda(), K.tile_4bit(p_b, s_b, 4096, 2048, qt, bs), 4096, 2048, qt, bs), x_b.float(), deq_b)
    assert torch.equal(y, K.qgemv_tiled(xd, tiled, N, Kdim, qt, bs)), "not deterministic / workspace not left clean"
    try:   # the k-split kernel on the same arguments: same operands, another summation order
        K.qgemv_tiled_tune(0, 37, True)
        y_v4 = K.qgemv_tiled(xd, tiled, N, Kdim, qt, bs)
    finally:
        K.qgemv_tiled_tune(0, 0, True)
    assert (y.float() - y_v4.float()).abs().max() <= 2 ** -6 * max(1e-3, y_v4.float().abs().max().item())
    # RESIDUAL epilogue in place + bias (model.py:591: x = attn(...) + x)
    bias, res = _randw((N,), 7, std=0.5), _randw((B, N), 8, std=1.0)
    buf = res.cuda().clone()
    out = K.qgemv_tiled(xd, tiled, N, Kdim, qt, bs, bias=bias.cuda(), residual=buf, out=buf)
    assert out.data_ptr() == buf.data_ptr()
    _check_gemv(out, x.float(), deq, post=lambda t: (t + bias.float()).to(torch.bfloat16).float() + res.float())
    K.check_gemv_errors(torch.device("cuda", 0))


@pytest.mark.parametrize("qt", ["nf4", "int4"])
@pytest.mark.parametrize("N,Kdim,B", [(13824, 5120, 16), (1376, 4096, 2), (344, 1024, 11)])
def test_qgemv_batched_tile_per_warp_swiglu(K, qt, N, Kdim, B):
    """The SwiGLU pair on the batched kernel: gate and up tiles alternate over the warps of a group, the up warp hands its
    rounded values to the gate warp through shared memory (model.py:712-716 rounding points)."""
    w1, x, p1, s1, bs, d1 = _gemv_case(K, N, Kdim, B, qt, seed=21)
    w2, _, p2, s2, _, d2 = _gemv_case(K, N, Kdim, B, qt, seed=22)
    t1, t2 = K.tile_4bit(p1, s1, N, Kdim, qt, bs), K.tile_4bit(p2, s2, N, Kdim, qt, bs)
    b1, b2 = _randw((N,), 23, std=0.5), _randw((N,), 24, std=0.5)
    for bias in (False, True):
        kw = dict(bias=b1.cuda(), bias2=b2.cuda()) if bias else {}
        y = K.qgemv_tiled(x.cuda(), t1, N, Kdim, qt, bs, tiled2=t2, **kw).float().cpu()
        xb = x.to(torch.bfloat16).cuda()
        a = torch.nn.functional.linear(xb, d1.to(torch.bfloat16).cuda(), b1.cuda() if bias else None)
        b = torch.nn.functional.linear(xb, d2.to(torch.bfloat16).cuda(), b2.cuda() if bias else None)
        comp = (torch.nn.functional.silu(a) * b).float().cpu()
        za, zb = x.double() @ d1.double().T + (b1.double() if bias else 0), x.double() @ d2.double().T + (b2.double() if bias else 0)
        wide = (torch.nn.functional.silu(za) * zb).float()
        e_k, e_t = (y - wide).abs().max().item(), (comp - wide).abs().max().item()
        assert e_k <= 2.0 * e_t + 2 ** -8 * wide.abs().max().item(), (bias, e_k, e_t)
        assert torch.equal(y, K.qgemv_tiled(x.cuda(), t1, N, Kdim, qt, bs, tiled2=t2, **kw).float().cpu())


@pytest.mark.parametrize("qt", ["nf4", "fp4"])
@pytest.mark.parametrize("N,Kdim,B,fold", [(4096, 1376, 1, 2), (256, 96, 3, 2), (512, 1376, 8, 2), (64, 48, 2, 4), (4096, 1376, 4, 2)])
def test_qgemv_tiled_folded_rows(K, qt, N, Kdim, B, fold):
    """A row-parallel shard whose K is not a multiple of the 64-element block (generate/tp.py:28-45: 11008 / 8 = 1376):
    bitsandbytes quantizes the shard's FLAT storage, so blocks straddle rows.  The same flat buffers viewed as
    (N / fold, fold * K) tile normally; the GEMV unfolds (b200_qgemv_tiled `fold`).  Checked against the oracle's
    dequantised (N, K) matrix, plain and with the residual epilogue."""
    w, x, packed, scale, bs, deq = _gemv_case(K, N, Kdim, B, qt, seed=N + B)
    assert K.tiled_bytes(N, Kdim, qt) == 0
    tiled = K.tile_4bit(packed, scale, N // fold, Kdim * fold, qt, bs)
    assert K.qgemv_tiled_supported(N, Kdim, qt, B, False, fold)
    y = K.qgemv_tiled(x.cuda(), tiled, N, Kdim, qt, bs, fold=fold)
    assert y.shape == (B, N)
    _check_gemv(y, x.float(), deq)
    res = _randw((B, N), 7, std=1.0)
    y2 = K.qgemv_tiled(x.cuda(), tiled, N, Kdim, qt, bs, fold=fold, residual=res.cuda())
    _check_gemv(y2, x.float(), deq, post=lambda t: t.to(torch.bfloat16).float() + res.float())
    # same numbers as the reference-layout kernel up to the order of the partial sums
    y_ref = K.qgemv(x.cuda(), packed, scale, N, Kdim, qt, bs)
    assert (y.float() - y_ref.float()).abs().max() <= 2 ** -6 * max(1e-3, y_ref.float().abs().max().item())


@pytest.mark.parametrize("qt", ["nf4", "int4"])
@pytest.mark.parametrize("N,Kdim,B", [(86, 32, 1), (33, 100, 2), (352, 96, 16), (7, 8, 1), (512, 1376, 1), (48, 1376, 5)])  # 1376 = 11008 / 8: TP-8 shard
def test_qgemv_generic_path_odd_shapes(K, qt, N, Kdim, B):
    w, x, packed, scale, bs, deq = _gemv_case(K, N, Kdim, B, qt, seed=N)
    y = K.qgemv(x.cuda(), packed, scale, N, Kdim, qt, bs)
    _check_gemv(y, x.float(), deq)


def test_qgemv_generic_equals_fast(K):
    w, x, packed, scale, bs, deq = _gemv_case(K, 512, 1024, 4, "nf4", seed=9)
    y_fast = K.qgemv(x.cuda(), packed, scale, 512, 1024, "nf4", bs)
    try:
        K.qgemv_tune(0, 0, force_generic=True)
        y_gen = K.qgemv(x.cuda(), packed, scale, 512, 1024, "nf4", bs)
    finally:
        K.qgemv_tune(0, 0, False)
    assert (y_fast.float() - y_gen.float()).abs().max() <= 2 ** -7 * y_gen.float().abs().max()


@pytest.mark.parametrize("Kdim", [4096, 11008, 96])
def test_qgemv_prologue_norm_and_residual_epilogue(K, Kdim):
    N, B = 512, 2
    w, x, packed, scale, bs, deq = _gemv_case(K, N, Kdim, B, "nf4", seed=Kdim)
    nw = (1.0 + 0.1 * _randw((Kdim,), 3, std=1.0).float()).to(torch.bfloat16)
    delta = _randw((B, Kdim), 4, std=0.5)
    res = _randw((B, N), 5, std=1.0)
    x_out = torch.empty(B, Kdim, dtype=torch.bfloat16, device="cuda")
    y = K.qgemv(x.cuda(), packed, scale, N, Kdim, "nf4", bs, delta=delta.cuda(), x_out=x_out, norm_weight=nw.cuda(),
                norm_eps=1e-5, residual=res.cuda())
    xs = (x + delta)  # bf16 add, as `x = h + x` in the reference (model.py:591)
    assert torch.equal(x_out.cpu(), xs)
    xn = M.rmsnorm(xs, nw, 1e-5)  # oracle rounding points (rmsnorm.py:19-25)
    h = (xn.double() @ deq.double().T).float().to(torch.bfloat16)
    want = (h + res).float()
    assert (y.float().cpu() - want).abs().max() <= 2 ** -6 * want.abs().max()


@pytest.mark.parametrize("N,Kdim,B", [(11008, 4096, 1), (352, 256, 3), (1376, 4096, 8)])
def test_qgemv_swiglu(K, N, Kdim, B):
    w1, x, p1, s1, bs, d1 = _gemv_case(K, N, Kdim, B, "nf4", seed=1)
    w2, _, p2, s2, _, d2 = _gemv_case(K, N, Kdim, B, "nf4", seed=2)
    y = K.qgemv(x.cuda(), p1, s1, N, Kdim, "nf4", bs, packed2=p2, scale2=s2)
    a = (x.double() @ d1.double().T).float().to(torch.bfloat16)
    b = (x.double() @ d2.double().T).float().to(torch.bfloat16)
    want = (torch.nn.functional.silu(a) * b).float()  # model.py:712-716 rounding points
    assert (y.float().cpu() - want).abs().max() <= 2 ** -6 * want.abs().max() + 1e-6


def test_qgemv_bias_and_errors(K):
    from lit_gpt_b200._cabi import B200Error

    w, x, packed, scale, bs, deq = _gemv_case(K, 256, 512, 1, "nf4", seed=3)
    bias = _randw((256,), 7, std=1.0)
    y = K.qgemv(x.cuda(), packed, scale, 256, 512, "nf4", bs, bias=bias.cuda())
    want = (x.double() @ deq.double().T).float() + bias.float()
    assert (y.float().cpu() - want).abs().max() <= 2 ** -6 * want.abs().max()
    with pytest.raises(B200Error, match="batch"):
        K.qgemv(torch.zeros(17, 512, dtype=torch.bfloat16, device="cuda"), packed, scale, 256, 512, "nf4", bs)
    with pytest.raises(B200Error, match="bfloat16"):
        K.qgemv(torch.zeros(1, 512, device="cuda"), packed, scale, 256, 512, "nf4", bs)
    with pytest.raises(B200Error, match="CUDA"):
        K.qgemv(torch.zeros(1, 512, dtype=torch.bfloat16), packed, scale, 256, 512, "nf4", bs)


def test_qgemv_linearity_full_size(K):
    """Size-independent property at the BASELINE shape (Llama-2-7B gate projection, 11008 x 4096):
    y(a*x1 + x2) tracks a*y(x1) + y(x2), and a one-hot x reads back a dequantised weight column exactly."""
    N, Kdim = 11008, 4096
    w = _randw((N, Kdim), 11)
    packed, absmax = K.quantize_4bit(w.cuda(), "nf4")
    onehot = torch.zeros(1, Kdim, dtype=torch.bfloat16, device="cuda")
    onehot[0, 1234] = 1.0
    col = K.qgemv(onehot, packed, absmax, N, Kdim)
    deq = K.dequantize_4bit(packed, absmax, (N, Kdim), torch.float32)
    # the GEMV multiplies bf16(code) by absmax in fp32 and rounds once: within 1 bf16 ulp of the exact dequant value
    assert (col[0].float() - deq[:, 1234]).abs().max() <= 2 ** -8 * deq[:, 1234].abs().max()
    x1, x2 = _randw((1, Kdim), 1, std=1.0).cuda(), _randw((1, Kdim), 2, std=1.0).cuda()
    lhs = K.qgemv((2 * x1 + x2), packed, absmax, N, Kdim).float()
    rhs = 2 * K.qgemv(x1, packed, absmax, N, Kdim).float() + K.qgemv(x2, packed, absmax, N, Kdim).float()
    assert (lhs - rhs).abs().max() <= 0.05 * rhs.abs().max()


# ---------------------------------------------------------------------------------------------------------
# norm / embedding / argmax
# ---------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("rows,dim", [(1, 4096), (5, 96), (16, 5120), (3, 8192)])
def test_rmsnorm(K, rows, dim):
    x, w = _randw((rows, dim), 1, std=1.0), (1 + _randw((dim,), 2, std=0.2).float()).to(torch.bfloat16)
    y = K.rmsnorm(x.cuda(), w.cuda(), 1e-5)
    want = M.rmsnorm(x, w, 1e-5)
    assert (y.cpu().float() - want.float()).abs().max() <= 2 ** -7 * want.float().abs().max()
    assert (y.cpu() == want).float().mean() > 0.99  # same rounding points: almost every element bit-equal
    d = _randw((rows, dim), 3, std=1.0)
    s = torch.empty_like(x).cuda()
    y2 = K.rmsnorm(x.cuda(), w.cuda(), 1e-5, delta=d.cuda(), sum_out=s)
    assert torch.equal(s.cpu(), x + d)
    assert (y2.cpu().float() - M.rmsnorm(x + d, w, 1e-5).float()).abs().max() <= 2 ** -7 * 8


def test_rmsnorm_golden_from_reference(K, golden):
    u = golden("units")
    x, w = torch.from_numpy(u["rms_x"]).bfloat16(), torch.from_numpy(u["rms_w"]).bfloat16()
    y = K.rmsnorm(x.cuda(), w.cuda(), 1e-5).cpu().float()
    ref = M.rmsnorm(x, w, 1e-5).float()
    assert (y - ref).abs().max() <= 2 ** -7 * ref.abs().max()


def test_embedding_and_argmax(K):
    wte = _randw((1000, 256), 1, std=1.0).cuda()
    idx = torch.tensor([[3], [999], [0]], device="cuda")
    assert torch.equal(K.embedding(idx, wte), wte[idx])
    logits = _randw((2, 32000), 5, std=1.0).cuda()
    logits[0, 777] = 50.0
    logits[0, 31999] = 50.0  # tie -> lowest index, as torch.argmax
    tok = torch.zeros(2, dtype=torch.int64, device="cuda")
    pos = torch.tensor([41], dtype=torch.int64, device="cuda")
    hist = torch.zeros(64, dtype=torch.int64, device="cuda")
    K.argmax_(logits, tok, pos_inout=pos, history=hist)
    assert tok.tolist() == [777, int(torch.argmax(logits[1]))] and pos.item() == 42 and hist[42].item() == 777


# ---------------------------------------------------------------------------------------------------------
# fused rope + kv append + decode attention
# ---------------------------------------------------------------------------------------------------------
def _attn_oracle(qkv, cos, sin, pos, kc, vc, H, G, hs, n_elem):
    """lit_gpt/model.py:624-665 restated with the oracle's ops on CPU (bf16 rope rounding, fp32 softmax)."""
    B = qkv.shape[0]
    qpk = H // G
    t = qkv.view(B, 1, G, qpk + 2, hs).permute(0, 2, 3, 1, 4)
    q, k, v = t.split((qpk, 1, 1), dim=2)
    q = q.reshape(B, H, 1, hs)
    k = k.reshape(B, G, 1, hs)
    v = v.reshape(B, G, 1, hs)
    c, s = cos[pos : pos + 1], sin[pos : pos + 1]
    q = torch.cat((M.apply_rope(q[..., :n_elem], c, s), q[..., n_elem:]), -1)
    k = torch.cat((M.apply_rope(k[..., :n_elem], c, s), k[..., n_elem:]), -1)
    kc, vc = kc.clone(), vc.clone()
    kc[:, :, pos] = k[:, :, 0]
    vc[:, :, pos] = v[:, :, 0]
    kk = kc[:, :, : pos + 1].float().repeat_interleave(qpk, dim=1)
    vv = vc[:, :, : pos + 1].float().repeat_interleave(qpk, dim=1)
    att = torch.softmax((q.float() @ kk.transpose(-1, -2)) / math.sqrt(hs), dim=-1)
    y = (att @ vv).transpose(1, 2).reshape(B, H * hs)
    return y, kc, vc


@pytest.mark.parametrize("H,G,hs", [(32, 32, 128), (32, 8, 128), (8, 1, 128), (8, 8, 64), (4, 2, 64), (12, 4, 64), (16, 1, 64)])
@pytest.mark.parametrize("pos,max_seq,B,splits", [(0, 64, 1, 0), (5, 64, 2, 1), (271, 272, 1, 0), (1000, 2048, 1, 7), (2047, 2048, 3, 0)])
@pytest.mark.parametrize("table_dtype", [torch.float32, torch.bfloat16])
def test_attn_decode(K, H, G, hs, pos, max_seq, B, splits, table_dtype):
    if table_dtype == torch.bfloat16 and (pos not in (5, 1000) or hs != 128):
        pytest.skip("bf16 tables: sampled subset")
    g = torch.Generator().manual_seed(H * 1000 + pos)
    n_elem = hs if H != 12 else hs // 4  # one partial-rotary case
    qkv = torch.randn(B, (H + 2 * G) * hs, generator=g).bfloat16()
    kc = torch.randn(B, G, max_seq, hs, generator=g).bfloat16()
    vc = torch.randn(B, G, max_seq, hs, generator=g).bfloat16()
    kc[:, :, pos:], vc[:, :, pos:] = 0, 0
    cos, sin = M.build_rope_cache(max_seq, n_elem, pos_dtype=torch.float32)
    cos, sin = cos.to(table_dtype), sin.to(table_dtype)
    want, kc_ref, vc_ref = _attn_oracle(qkv, cos, sin, pos, kc, vc, H, G, hs, n_elem)
    kcd, vcd = kc.cuda(), vc.cuda()
    y = K.attn_decode(qkv.cuda(), cos.cuda(), sin.cuda(), torch.tensor([pos], device="cuda"), kcd, vcd, H, G, hs, n_elem,
                      num_splits=splits)
    assert torch.equal(kcd.cpu(), kc_ref) and torch.equal(vcd.cpu(), vc_ref), "cache append differs from the oracle"
    assert (y.cpu().float() - want).abs().max() <= 2 ** -7 * max(1.0, want.abs().max().item())
    # preselected table row (module path) gives the same answer
    y2 = K.attn_decode(qkv.cuda(), cos[pos : pos + 1].cuda(), sin[pos : pos + 1].cuda(), torch.tensor([pos], device="cuda", dtype=torch.int32),
                       kc.cuda(), vc.cuda(), H, G, hs, n_elem, num_splits=splits, table_preselected=True)
    assert torch.equal(y, y2)


@pytest.mark.parametrize("H,G,n_elem", [(64, 8, 128), (32, 4, 128), (16, 1, 128), (6, 2, 128), (8, 2, 32), (32, 16, 128)])
@pytest.mark.parametrize("pos,max_seq,B,splits", [(0, 40, 1, 0), (31, 64, 2, 1), (32, 64, 1, 2), (700, 1024, 2, 0), (4095, 4096, 1, 0), (130, 300, 1, 64)])
def test_attn_decode_gqa_mma_vs_oracle_and_scalar(K, H, G, n_elem, pos, max_seq, B, splits):
    """Grouped-query decode attention on mma.sync (head size 128, q_per_kv 2..16) against the oracle formulation and against the
    scalar kernel on the same inputs (same cache append bit for bit; outputs within the bf16 bar of test_attn_decode)."""
    hs = 128
    g = torch.Generator().manual_seed(H * 77 + pos)
    qkv = torch.randn(B, (H + 2 * G) * hs, generator=g).bfloat16()
    kc = torch.randn(B, G, max_seq, hs, generator=g).bfloat16()
    vc = torch.randn(B, G, max_seq, hs, generator=g).bfloat16()
    kc[:, :, pos:], vc[:, :, pos:] = 1e4, 1e4   # stale rows past the filled region must never be read
    cos, sin = M.build_rope_cache(max_seq, n_elem, pos_dtype=torch.float32)
    want, kc_ref, vc_ref = _attn_oracle(qkv, cos, sin, pos, kc, vc, H, G, hs, n_elem)
    posd = torch.tensor([pos], device="cuda")
    kcd, vcd = kc.cuda(), vc.cuda()
    y = K.attn_decode(qkv.cuda(), cos.cuda(), sin.cuda(), posd, kcd, vcd, H, G, hs, n_elem, num_splits=splits)
    assert torch.equal(kcd.cpu(), kc_ref) and torch.equal(vcd.cpu(), vc_ref), "cache append differs from the oracle"
    tol = 2 ** -7 * max(1.0, want.abs().max().item())
    assert (y.cpu().float() - want).abs().max() <= tol
    prev = K.attn_decode_tune(force_scalar=True)
    try:
        kcs, vcs = kc.cuda(), vc.cuda()
        ys = K.attn_decode(qkv.cuda(), cos.cuda(), sin.cuda(), posd, kcs, vcs, H, G, hs, n_elem, num_splits=splits)
    finally:
        K.attn_decode_tune(force_scalar=prev)
    assert torch.equal(kcs, kcd) and torch.equal(vcs, vcd)
    assert (y.float() - ys.float()).abs().max().item() <= tol


def test_rope_kv_append(K):
    B, T, H, G, hs, max_seq = 2, 9, 8, 2, 128, 32
    g = torch.Generator().manual_seed(0)
    qkv = torch.randn(B, T, (H + 2 * G) * hs, generator=g).bfloat16()
    cos, sin = M.build_rope_cache(max_seq, hs, pos_dtype=torch.float32)
    pos = torch.arange(3, 3 + T)
    kc = torch.zeros(B, G, max_seq, hs, dtype=torch.bfloat16, device="cuda")
    vc = torch.zeros_like(kc)
    q = K.rope_kv_append(qkv.cuda(), cos.cuda(), sin.cuda(), pos.cuda(), kc, vc, H, G, hs, hs)
    qpk = H // G
    t = qkv.view(B, T, G, qpk + 2, hs).permute(0, 2, 3, 1, 4)
    qq, kk, vv = t.split((qpk, 1, 1), dim=2)
    q_ref = M.apply_rope(qq.reshape(B, H, T, hs), cos[pos], sin[pos])
    k_ref = M.apply_rope(kk.reshape(B, G, T, hs), cos[pos], sin[pos])
    assert torch.equal(q.cpu(), q_ref)
    assert torch.equal(kc[:, :, 3 : 3 + T].cpu(), k_ref) and torch.equal(vc[:, :, 3 : 3 + T].cpu(), vv.reshape(B, G, T, hs))
    assert kc[:, :, :3].abs().sum() == 0 and kc[:, :, 3 + T :].abs().sum() == 0


@pytest.mark.parametrize("B,T,H,G,hs,start,max_seq", [
    (1, 16, 8, 8, 128, 0, 64),        # MHA, one partial tile
    (2, 200, 8, 2, 128, 0, 256),      # GQA, several key tiles, ragged last query tile
    (1, 130, 4, 1, 64, 37, 300),      # MQA, head size 64 (pythia), chunked prefill starting at position 37
    (1, 1, 4, 4, 128, 5, 16),         # single row through the prefill kernel
    (1, 513, 2, 2, 128, 0, 513),      # cache exactly full
])
def test_attn_prefill_vs_oracle(K, B, T, H, G, hs, start, max_seq):
    """Prefill attention against the reference formulation (lit_gpt/model.py:658-665: SDPA with the mask rows that
    input_pos selects), computed in fp32 on the bf16 inputs.  Tolerance: bf16 output rounding (2^-8 relative) plus
    the bf16 rounding of the probabilities fed to P.V, i.e. 2^-7 * max|v| absolute."""
    g = torch.Generator().manual_seed(B * 1000 + T)
    q = torch.randn(B, H, T, hs, generator=g).bfloat16()
    kc = torch.zeros(B, G, max_seq, hs).bfloat16()
    vc = torch.zeros(B, G, max_seq, hs).bfloat16()
    filled = start + T
    kc[:, :, :filled] = torch.randn(B, G, filled, hs, generator=g).bfloat16()
    vc[:, :, :filled] = torch.randn(B, G, filled, hs, generator=g).bfloat16()
    # garbage past the filled region must never be read (reference: masked out)
    kc[:, :, filled:] = 1e4
    vc[:, :, filled:] = 1e4
    pos = torch.arange(start, start + T)
    y = K.attn_prefill(q.cuda(), kc.cuda(), vc.cuda(), pos.cuda(), 1.0 / hs ** 0.5).float().cpu()
    y32 = K.attn_prefill(q.cuda(), kc.cuda(), vc.cuda(), pos.to(torch.int32).cuda(), 1.0 / hs ** 0.5).float().cpu()
    assert torch.equal(y, y32)
    qpk = H // G
    kf = kc.float().repeat_interleave(qpk, dim=1)
    vf = vc.float().repeat_interleave(qpk, dim=1)
    mask = torch.tril(torch.ones(max_seq, max_seq, dtype=torch.bool))[pos].view(1, 1, T, max_seq)   # model.py:509, 802-804
    ref = torch.nn.functional.scaled_dot_product_attention(q.float(), kf, vf, attn_mask=mask, dropout_p=0.0, scale=1.0 / hs ** 0.5)
    ref = ref.transpose(1, 2).reshape(B, T, H * hs)
    err = (y - ref).abs().max().item()
    tol = 2.0 ** -7 * vc[:, :, :filled].float().abs().max().item()
    assert err <= tol, (err, tol)


# ---------------------------------------------------------------------------------------------------------
# tiled layout + streaming GEMV (the decode hot path)
# ---------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("qt", ["nf4", "fp4", "int4"])
@pytest.mark.parametrize("N,Kdim", [(512, 4096), (344, 11008), (1000, 5120), (20, 128), (7, 256), (4096, 1024)])
def test_tile_untile_roundtrip_bit_exact(K, qt, N, Kdim):
    """The tiled layout is a pure permutation of the reference bytes: untile(tile(.)) is the identity, bit for bit."""
    w, x, packed, scale, bs, deq = _gemv_case(K, N, Kdim, 1, qt, seed=N)
    tiled = K.tile_4bit(packed, scale, N, Kdim, qt, bs)
    assert tiled.numel() == K.tiled_bytes(N, Kdim, qt) > 0
    p2, s2 = K.untile_4bit(tiled, N, Kdim, qt, bs)
    assert torch.equal(p2.view(torch.uint8).flatten(), packed.view(torch.uint8).flatten())
    assert torch.equal(s2.view(torch.uint8).flatten(), scale.view(torch.uint8).flatten())


def test_tiled_layout_unsupported_shapes(K):
    assert K.tiled_bytes(64, 96, "nf4") == 0 and K.tiled_bytes(64, 64, "int4") == 0 and K.tiled_bytes(64, 64, "nf4") > 0


@pytest.mark.parametrize("qt", ["nf4", "fp4", "int4"])
@pytest.mark.parametrize("N,Kdim,B", [(512, 4096, 1), (4096, 4096, 1), (344, 11008, 1), (1000, 5120, 3), (4096, 1024, 8),
                                      (2048, 1024, 9), (256, 256, 5), (20, 128, 1), (512, 8192, 2), (12288, 4096, 1), (40, 1024, 16)])
@pytest.mark.parametrize("depth", [0, 2, 4])
def test_qgemv_tiled(K, qt, N, Kdim, B, depth):
    if depth and (N, Kdim, B) not in ((512, 4096, 1), (344, 11008, 1), (1000, 5120, 3)):
        pytest.skip("depth variants: sampled subset")
    w, x, packed, scale, bs, deq = _gemv_case(K, N, Kdim, B, qt, seed=N + B)
    tiled = K.tile_4bit(packed, scale, N, Kdim, qt, bs)
    try:
        K.qgemv_tiled_tune(depth, 0, True)
        y = K.qgemv_tiled(x.cuda(), tiled, N, Kdim, qt, bs, prefetch=tiled)
        y2 = K.qgemv_tiled(x.cuda(), tiled, N, Kdim, qt, bs)
    finally:
        K.qgemv_tiled_tune(0, 0, True)
    assert y.shape == (B, N) and y.dtype == torch.bfloat16
    _check_gemv(y, x.float(), deq)
    assert torch.equal(y, y2), "not deterministic"
    y_ref = K.qgemv(x.cuda(), packed, scale, N, Kdim, qt, bs)  # reference-layout kernel: same arithmetic, other k order
    assert (y.float() - y_ref.float()).abs().max() <= 2 ** -6 * max(1e-3, y_ref.float().abs().max().item())


@pytest.mark.parametrize("qt", ["nf4", "int4"])
@pytest.mark.parametrize("Kdim", [4096, 11008, 128])
def test_qgemv_tiled_prologue_norm_and_residual_epilogue(K, qt, Kdim):
    N, B = 520, 2
    w, x, packed, scale, bs, deq = _gemv_case(K, N, Kdim, B, qt, seed=Kdim)
    tiled = K.tile_4bit(packed, scale, N, Kdim, qt, bs)
    nw = (1.0 + 0.1 * _randw((Kdim,), 3, std=1.0).float()).to(torch.bfloat16)
    delta = _randw((B, Kdim), 4, std=0.5)
    res = _randw((B, N), 5, std=1.0)
    x_out = torch.empty(B, Kdim, dtype=torch.bfloat16, device="cuda")
    y = K.qgemv_tiled(x.cuda(), tiled, N, Kdim, qt, bs, delta=delta.cuda(), x_out=x_out, norm_weight=nw.cuda(), norm_eps=1e-5,
                      residual=res.cuda())
    xs = (x + delta)  # bf16 add, as `x = h + x` in the reference (model.py:591)
    assert torch.equal(x_out.cpu(), xs)
    xn = M.rmsnorm(xs, nw, 1e-5)  # oracle rounding points (rmsnorm.py:19-25)
    h = (xn.double() @ deq.double().T).float().to(torch.bfloat16)
    want = (h + res).float()
    assert (y.float().cpu() - want).abs().max() <= 2 ** -6 * want.abs().max()


@pytest.mark.parametrize("qt", ["nf4", "int4"])
@pytest.mark.parametrize("N,Kdim,B", [(11008, 4096, 1), (352, 256, 3), (1376, 4096, 8), (1000, 1024, 12)])
def test_qgemv_tiled_swiglu(K, qt, N, Kdim, B):
    w1, x, p1, s1, bs, d1 = _gemv_case(K, N, Kdim, B, qt, seed=1)
    w2, _, p2, s2, _, d2 = _gemv_case(K, N, Kdim, B, qt, seed=2)
    t1, t2 = K.tile_4bit(p1, s1, N, Kdim, qt, bs), K.tile_4bit(p2, s2, N, Kdim, qt, bs)
    y = K.qgemv_tiled(x.cuda(), t1, N, Kdim, qt, bs, tiled2=t2)
    a = (x.double() @ d1.double().T).float().to(torch.bfloat16)
    b = (x.double() @ d2.double().T).float().to(torch.bfloat16)
    want = (torch.nn.functional.silu(a) * b).float()  # model.py:712-716 rounding points
    assert (y.float().cpu() - want).abs().max() <= 2 ** -6 * want.abs().max() + 1e-6


def test_qgemv_tiled_bias_and_errors(K):
    from lit_gpt_b200._cabi import B200Error

    w, x, packed, scale, bs, deq = _gemv_case(K, 256, 512, 1, "nf4", seed=3)
    tiled = K.tile_4bit(packed, scale, 256, 512, "nf4", bs)
    bias = _randw((256,), 7, std=1.0)
    y = K.qgemv_tiled(x.cuda(), tiled, 256, 512, "nf4", bs, bias=bias.cuda())
    want = (x.double() @ deq.double().T).float() + bias.float()
    assert (y.float().cpu() - want).abs().max() <= 2 ** -6 * want.abs().max()
    with pytest.raises(B200Error, match="batch"):
        K.qgemv_tiled(torch.zeros(17, 512, dtype=torch.bfloat16, device="cuda"), tiled, 256, 512, "nf4", bs)
    # activation rows that do not fit beside the ring run through k-windows now (test_qgemv_tiled_k_windows); the tensor-
    # parallel consumer is the one remaining shape outside every plan and must say so without launching anything
    assert K.qgemv_tiled_supported(16, 8192, "nf4", 16) and not K.qgemv_tiled_supported(16, 8192, "nf4", 17)


def test_qgemv_tiled_onehot_reads_dequantised_column(K):
    """Size-independent property at the BASELINE shape (Llama-2-7B gate projection 11008 x 4096)."""
    N, Kdim = 11008, 4096
    w = _randw((N, Kdim), 11)
    packed, absmax = K.quantize_4bit(w.cuda(), "nf4")
    tiled = K.tile_4bit(packed, absmax, N, Kdim)
    onehot = torch.zeros(1, Kdim, dtype=torch.bfloat16, device="cuda")
    onehot[0, 1234] = 1.0
    col = K.qgemv_tiled(onehot, tiled, N, Kdim)
    deq = K.dequantize_4bit(packed, absmax, (N, Kdim), torch.float32)
    assert (col[0].float() - deq[:, 1234]).abs().max() <= 2 ** -8 * deq[:, 1234].abs().max()


# ---- prefill GEMM on tcgen05 / TMEM (csrc/qgemm.cu) ---------------------------------------------------------------
@pytest.mark.parametrize("qt", ["nf4", "fp4", "int4"])
@pytest.mark.parametrize("M,N,Kdim", [(128, 128, 128), (17, 8, 128), (256, 256, 512), (300, 520, 1024), (129, 4096, 4096), (1000, 344, 11008),
                                      (64, 4096, 4096), (128, 4096, 11008), (17, 1000, 5120), (100, 12288, 4096), (200, 4096, 4096), (520, 2048, 4096), (1153, 1024, 2048)])
def test_qgemm_tcgen05(K, qt, M, N, Kdim):
    w, _, packed, scale, bs, deq = _gemv_case(K, N, Kdim, 1, qt, seed=M + N)
    x = _randw((M, Kdim), 5, std=1.0)
    bias = _randw((N,), 6, std=0.5) if M in (300, 100, 17) else None
    y = K.qgemm(x.cuda(), packed, scale, N, Kdim, qt, bs, bias=None if bias is None else bias.cuda())
    assert y.shape == (M, N)
    from lit_gpt_b200 import _cabi as C
    need = C.load().b200_qgemm_workspace_bytes(M, N, Kdim)   # short prompts are split over k ranges: fp32 partial sums, summed in range order
    if need:
        assert need >= 2 * M * N * 4
        assert torch.equal(y, K.qgemm(x.cuda(), packed, scale, N, Kdim, qt, bs, bias=None if bias is None else bias.cuda())), "not deterministic"
    if M <= 256 and Kdim >= 512 and N <= 128 * 74:
        assert need > 0, "output tiles fill less than half the SMs: the launch must be split over k ranges"
    # the weight tile the tensor core sees is bf16_rn(dequant): same operands as the reference's dequantize + F.linear,
    # so the only difference to a library GEMM on those operands is the fp32 accumulation order
    wb = deq.to(torch.bfloat16)
    ref = torch.nn.functional.linear(x.cuda(), wb.cuda(), None if bias is None else bias.cuda()).float().cpu()
    wide = x.double() @ wb.double().T + (0 if bias is None else bias.double())
    e_k = (y.float().cpu().double() - wide).abs().max().item()
    e_t = (ref.double() - wide).abs().max().item()
    scale_y = wide.abs().max().item()
    assert e_k <= 2.0 * e_t + 2 ** -8 * scale_y, f"kernel err {e_k:.4g} vs library-GEMM err {e_t:.4g} (|y|max {scale_y:.3g})"


def test_qgemm_onehot_rows_read_dequantised_weights(K):
    """x = identity block: y must be EXACTLY bf16(dequant(W)) -- checks the swizzled operand layout, the code book,
    the nibble order and the TMEM read-back mapping without any accumulation error."""
    N, Kdim = 136, 256
    for qt in ("nf4", "int4"):
        w, _, packed, scale, bs, deq = _gemv_case(K, N, Kdim, 1, qt, seed=3)
        x = torch.eye(Kdim, dtype=torch.bfloat16)
        y = K.qgemm(x.cuda(), packed, scale, N, Kdim, qt, bs)
        assert torch.equal(y.cpu(), deq.to(torch.bfloat16).T.contiguous()), qt


def test_qgemm_unsupported_shapes(K):
    w, _, packed, scale, bs, _ = _gemv_case(K, 20, 96, 1, "nf4")
    with pytest.raises(NotImplementedError):
        K.qgemm(torch.zeros(32, 96, dtype=torch.bfloat16, device="cuda"), packed, scale, 20, 96, "nf4", 64)


# ---- double quantisation of absmax (bnb.nf4-dq / fp4-dq) ---------------------------------------------------------
def test_double_quant_absmax_bit_exact(K):
    g = torch.Generator().manual_seed(11)
    for n in (256, 1000, 4096 * 64, 5):
        a = (torch.rand(n, generator=g) * 0.08 + 0.01).float()
        if n == 1000:
            a[300:556] = 0.05  # a block whose centred values are all equal
        q, a2, off, deq = K.double_quant_absmax(a.cuda())
        q_ref, a2_ref, off_ref = Q.double_quantize_absmax(a.numpy(), offset=off)
        assert np.array_equal(q.cpu().numpy(), q_ref) and np.array_equal(a2.cpu().numpy(), a2_ref)
        assert np.array_equal(deq.cpu().numpy(), Q.dequantize_absmax(q_ref, a2_ref, off_ref))
        assert abs(off - float(a.mean())) <= 1e-6
    assert np.array_equal(K.dynamic_map(torch.device("cuda")).cpu().numpy(), Q.DYNAMIC_MAP)


@pytest.mark.parametrize("mode", ["nf4-dq", "fp4-dq"])
def test_qlinear_double_quant_modes(K, mode):
    """--quantize bnb.nf4-dq: weights dequantise to code * dequant8(absmax) exactly as the oracle says, and the GEMV /
    prefill GEMM of such a layer agree with the dequantised matrix."""
    from lit_gpt_b200.quant import QLinear

    qt = mode.split("-")[0]
    w = _randw((512, 1024), 21)
    lin = QLinear(1024, 512, bias=False, mode=mode)
    lin.weight.data = w.clone()
    lin = lin.cuda()
    assert lin.is_quantized and lin.quant_state[0].dtype == torch.uint8 and lin.quant_state[4] is not None
    packed, absmax = QC.quantize_4bit(w.float().numpy(), qt)
    qa, a2, off = Q.double_quantize_absmax(absmax, offset=lin.quant_state[4][0])
    deq = QC.dequantize_4bit(packed, Q.dequantize_absmax(qa, a2, off), (512, 1024), qt)
    assert torch.equal(lin.dequantize(torch.float32).cpu(), torch.from_numpy(deq))
    x = _randw((1, 1024), 22, std=1.0)
    _check_gemv(lin(x.cuda()).cpu(), x, torch.from_numpy(deq))
    xm = _randw((64, 1024), 23, std=1.0)
    _check_gemv(lin(xm.cuda()).cpu(), xm, torch.from_numpy(deq))
