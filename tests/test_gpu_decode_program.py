"""Persistent decode-step kernel (csrc/decode_step.cu) against the per-op kernels it fuses and against the oracle.

The per-op kernels are themselves checked against the oracle in test_gpu_kernels.py; here the same ops are run as
ONE program (GEMV / ATTN / EMBED with grid barriers in between) and must reproduce them bit for bit: same inner
loop, same reduction order, same rounding points.
"""
import numpy as np
import pytest
import torch

from oracle import model_ref as M
from test_gpu_kernels import _attn_oracle, _check_gemv, _gemv_case

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def K():
    from lit_gpt_b200 import kernels

    return kernels


def _tiled(K, N, Kdim, B, qt, seed):
    w, x, p, s, bs, deq = _gemv_case(K, N, Kdim, B, qt, seed=seed)
    return K.tile_4bit(p, s, N, Kdim, qt, bs), x, bs, deq


@pytest.mark.parametrize("qt", ["nf4", "fp4", "int4"])
@pytest.mark.parametrize("B", [1, 3, 12])
@pytest.mark.parametrize("lookahead", [0, 256])
def test_program_gemv_chain_matches_per_op_kernels(K, qt, B, lookahead):
    d, I = 1024, (2816 if B < 12 else 1408)  # batch 12 x K must fit the shared-memory plan
    t_up, x, bs, deq_up = _tiled(K, I, d, B, qt, 1)
    t_gate, _, _, _ = _tiled(K, I, d, B, qt, 2)
    t_down, _, _, deq_down = _tiled(K, d, I, B, qt, 3)
    t_head, _, _, _ = _tiled(K, 1000, d, B, qt, 4)
    xd = x.cuda()
    nw = (torch.randn(d) * 0.1 + 1).bfloat16().cuda()
    # per-op kernels
    act_ref = K.qgemv_tiled(xd, t_gate, I, d, qt, bs, tiled2=t_up, norm_weight=nw, norm_eps=1e-5)
    x2_ref = K.qgemv_tiled(act_ref, t_down, d, I, qt, bs, residual=xd)
    head_ref = K.qgemv_tiled(x2_ref, t_head, 1000, d, qt, bs, norm_weight=nw, norm_eps=1e-5)
    # one program
    K.decode_step_tune(lookahead)
    try:
        P = K.DecodeProgram(B, qt, bs, xd.device)
        act, x2, head = (torch.zeros(B, n, dtype=torch.bfloat16, device="cuda") for n in (I, d, 1000))
        P.gemv(xd, t_gate, I, d, act, tiled2=t_up, norm_weight=nw, norm_eps=1e-5)
        P.gemv(act, t_down, d, I, x2, residual=xd)
        P.gemv(x2, t_head, 1000, d, head, norm_weight=nw, norm_eps=1e-5)
        P.finalize()
        for _ in range(3):  # replays reuse the barrier counter and the ring state
            act.zero_(); x2.zero_(); head.zero_()
            P.run()
            torch.cuda.synchronize()
            assert torch.equal(act, act_ref) and torch.equal(x2, x2_ref) and torch.equal(head, head_ref)
    finally:
        K.decode_step_tune(0)


def test_program_gemv_delta_x_out_and_small_n(K):
    # N smaller than one tile per CTA (most CTAs idle), delta + x_out publication, bias
    B, d, N = 2, 512, 40
    t, x, bs, deq = _tiled(K, N, d, B, "nf4", 7)
    xd = x.cuda()
    delta = (torch.randn(B, d) * 0.5).bfloat16().cuda()
    bias = torch.randn(N).bfloat16().cuda()
    x_out_ref = torch.empty_like(xd)
    y_ref = K.qgemv_tiled(xd, t, N, d, "nf4", bs, delta=delta, x_out=x_out_ref, bias=bias)
    P = K.DecodeProgram(B, "nf4", bs, xd.device)
    y, x_out = torch.zeros(B, N, dtype=torch.bfloat16, device="cuda"), torch.zeros_like(xd)
    P.gemv(xd, t, N, d, y, delta=delta, x_out=x_out, bias=bias)
    P.finalize().run()
    torch.cuda.synchronize()
    assert torch.equal(y, y_ref) and torch.equal(x_out, x_out_ref)
    _check_gemv(y.cpu(), (xd + delta).cpu(), deq, post=lambda t: t + bias.cpu().float())


@pytest.mark.parametrize("H,G,hs", [(32, 32, 128), (32, 8, 128), (8, 1, 128), (8, 8, 64), (12, 4, 64)])
@pytest.mark.parametrize("pos,max_seq,B,splits", [(0, 64, 1, 0), (271, 272, 1, 0), (1000, 2048, 2, 7), (2047, 2048, 3, 0)])
def test_program_attn_matches_kernel_and_oracle(K, H, G, hs, pos, max_seq, B, splits):
    g = torch.Generator().manual_seed(H * 1000 + pos)
    n_elem = hs if H != 12 else hs // 4
    qkv = torch.randn(B, (H + 2 * G) * hs, generator=g).bfloat16()
    kc = torch.randn(B, G, max_seq, hs, generator=g).bfloat16()
    vc = torch.randn(B, G, max_seq, hs, generator=g).bfloat16()
    kc[:, :, pos:], vc[:, :, pos:] = 0, 0
    cos, sin = M.build_rope_cache(max_seq, n_elem, pos_dtype=torch.float32)
    want, kc_ref, vc_ref = _attn_oracle(qkv, cos, sin, pos, kc, vc, H, G, hs, n_elem)
    posd = torch.tensor([pos], device="cuda")
    kcd, vcd = kc.cuda(), vc.cuda()
    # a program needs at least one GEMV op: project the attention output with a small matrix afterwards
    d = H * hs
    t, _, bs, _ = _tiled(K, 64, d, B, "nf4", 5)
    y = torch.zeros(B, d, dtype=torch.bfloat16, device="cuda")
    z = torch.zeros(B, 64, dtype=torch.bfloat16, device="cuda")
    P = K.DecodeProgram(B, "nf4", bs, y.device)
    P.attn(qkv.cuda(), cos.cuda(), sin.cuda(), posd, kcd, vcd, y, H, G, hs, n_elem, num_splits=splits)
    P.gemv(y, t, 64, d, z)
    P.finalize().run()
    torch.cuda.synchronize()
    assert torch.equal(kcd.cpu(), kc_ref) and torch.equal(vcd.cpu(), vc_ref), "cache append differs from the oracle"
    assert (y.cpu().float() - want).abs().max() <= 2 ** -7 * max(1.0, want.abs().max().item())
    assert torch.equal(z, K.qgemv_tiled(y, t, 64, d, "nf4", bs))
    if splits:  # same split count => same merge order => identical bits as the stand-alone kernel
        prev = K.attn_decode_tune(force_scalar=True)  # the program runs the scalar body for every shape
        try:
            y_k = K.attn_decode(qkv.cuda(), cos.cuda(), sin.cuda(), posd, kc.cuda(), vc.cuda(), H, G, hs, n_elem, num_splits=splits)
        finally:
            K.attn_decode_tune(force_scalar=prev)
        assert torch.equal(y, y_k)


def test_program_embed_and_errors(K):
    B, d, V = 3, 256, 100
    wte = torch.randn(V, d).bfloat16().cuda()
    idx = torch.tensor([5, 99, 0], device="cuda")
    t, _, bs, _ = _tiled(K, 32, d, B, "nf4", 9)
    x, y = torch.zeros(B, d, dtype=torch.bfloat16, device="cuda"), torch.zeros(B, 32, dtype=torch.bfloat16, device="cuda")
    P = K.DecodeProgram(B, "nf4", bs, x.device)
    P.embed(idx, wte, x)
    P.gemv(x, t, 32, d, y)
    P.finalize().run()
    torch.cuda.synchronize()
    assert torch.equal(x, wte[idx]) and torch.equal(y, K.qgemv_tiled(wte[idx], t, 32, d, "nf4", bs))
    # validation
    from lit_gpt_b200._cabi import B200Error

    with pytest.raises(B200Error, match="alias"):
        Q = K.DecodeProgram(B, "nf4", bs, x.device)
        Q.gemv(x, t, 32, d, x[:, :32].contiguous() if False else y, residual=y)
        Q.finalize()
    with pytest.raises(B200Error, match="at least one GEMV"):
        Q = K.DecodeProgram(B, "nf4", bs, x.device)
        Q.embed(idx, wte, x)
        Q.finalize()
    with pytest.raises(NotImplementedError, match="shared memory"):
        Q = K.DecodeProgram(16, "nf4", bs, x.device)
        tb = torch.zeros(K.tiled_bytes(16, 8192, "nf4"), dtype=torch.uint8, device="cuda")
        Q.gemv(torch.zeros(16, 8192, dtype=torch.bfloat16, device="cuda"), tb, 16, 8192, torch.zeros(16, 16, dtype=torch.bfloat16, device="cuda"))
        Q.finalize()


@pytest.mark.parametrize("mode", ["nf4", "int4"])
@torch.inference_mode()
def test_model_program_equals_chain(mode):
    """GPT(T=1) through ONE persistent kernel == the same step as a chain of per-op launches, bit for bit, over a
    rollout that appends to the KV cache (Llama-2-7B widths, 2 blocks)."""
    from lit_gpt_b200.config import name_to_config
    from test_gpu_model import build_cuda_model

    kw = dict(name_to_config["Llama-2-7b-hf"], n_layer=2)
    from lit_gpt_b200 import Config

    rc = M.RefConfig.from_any(Config(**kw))
    p = M.init_weights(rc, seed=4321, dtype=torch.bfloat16)
    model = build_cuda_model(kw, p, mode, max_seq=24)
    prompt = torch.randint(0, 32000, (1, 6), generator=torch.Generator().manual_seed(1)).cuda()
    outs = {}
    for use in (False, True):
        model.use_decode_program = use
        for blk in model.transformer.h:
            blk.attn.kv_cache.reset_parameters()
        model(prompt, torch.arange(6, device="cuda"))
        tok, rows = torch.tensor([[11]], device="cuda"), []
        for i in range(6, 14):
            logits = model(tok, torch.tensor([i], device="cuda"))
            rows.append(logits[0, -1].clone())
            tok = logits[0, -1].argmax().view(1, 1)
        outs[use] = torch.stack(rows)
        if use:
            assert model.__dict__.get("_decode_programs"), "the persistent kernel did not run"
    assert torch.equal(outs[False], outs[True])
