"""End-to-end parity of the quantized CUDA path (GPT / generate through the public API) against the oracle:
the reference's model restated in oracle/model_ref.py (pinned on reference goldens) with every Linear weight replaced
by dequant(quant(W)) from oracle/quant_ref.py — exactly what `--quantize bnb.nf4` computes.

Tolerance (stated, as north_star asks): max-abs logit error of the CUDA path against the fp32 oracle must be
<= 2x the error of the reference formulation run in bf16 (oracle in bf16 on CPU) + 1e-3 * max|logit|; greedy tokens
must agree wherever the oracle's top-1/top-2 margin exceeds that tolerance.
"""
import ast

import numpy as np
import pytest
import torch

from oracle import model_ref as M
from oracle import quant_ref as Q
from oracle import quant_ref_c as QC

pytestmark = pytest.mark.gpu


def _linear_keys(p):
    return [k for k in p if k.endswith(".weight") and p[k].dim() == 2 and "wte" not in k]


def oracle_weights(p, mode):
    """fp32 state dict whose Linear weights are dequant(quant(W_bf16)) — the oracle's quantized model."""
    out = {}
    for k, v in p.items():
        v16 = v.to(torch.bfloat16)
        if k in _linear_keys(p):
            w = v16.float().numpy()
            if mode == "int4":
                qw, s, z = Q.quantize_int4_g128(w)
                out[k] = torch.from_numpy(QC.dequantize_int4(qw, s, z, w.shape[1]))
            else:
                pk, am = QC.quantize_4bit(w, mode)
                out[k] = torch.from_numpy(QC.dequantize_4bit(pk, am, w.shape, mode))
        else:
            out[k] = v16.float()
    return out


def build_cuda_model(cfg_kwargs, p, mode, max_seq, batch=1):
    from lit_gpt_b200 import GPT, Config, QuantPrecision
    from lit_gpt_b200.fabric import Fabric

    fabric = Fabric(devices=1, plugins=QuantPrecision(mode, torch.bfloat16))
    with fabric.init_module(empty_init=True):
        model = GPT(Config(**cfg_kwargs))
    with fabric.init_tensor():
        model.max_seq_length = max_seq
        model.set_kv_cache(batch_size=batch)
    model.eval()
    model = fabric.setup_module(model)
    model.load_state_dict({k: v.to(torch.bfloat16) for k, v in p.items()}, strict=True)  # quantizes on load
    return model


def cfg_from_golden(g):
    cfgd = ast.literal_eval(str(g["config"]))
    kw = {k: v for k, v in cfgd.items() if k not in ("head_size", "rope_n_elem")}
    kw["rotary_percentage"] = cfgd["rope_n_elem"] / cfgd["head_size"]
    return cfgd, kw


def run_oracle(rc, pw, cos, sin, prompt, n_total, dtype, forced=None):
    pw = {k: v.to(dtype) for k, v in pw.items()}
    st = M.RefState(rc, n_total, 1, dtype, cos=cos, sin=sin)
    toks, logits = M.generate(rc, pw, st, prompt, n_total, temperature=0.0, return_logits=True, forced=forced)
    return toks, logits.float()


def check_rollout(model, rc, pw, prompt, n_total, label):
    import lit_gpt_b200.generate.base as gb

    cos, sin = model.cos.cpu(), model.sin.cpu()
    toks32, logits32 = run_oracle(rc, pw, cos.float(), sin.float(), prompt, n_total, torch.float32)
    # the bf16 reference formulation follows the fp32 oracle's tokens, so that its error is a per-step rounding error and
    # not the distance between two diverged rollouts (which would inflate the tolerance)
    _, logits16 = run_oracle(rc, pw, cos, sin, prompt, n_total, torch.bfloat16, forced=toks32)
    # teacher-forced logits of the CUDA path along the ORACLE's token sequence (so errors do not compound)
    T = prompt.numel()
    for blk in model.transformer.h:
        blk.attn.kv_cache.reset_parameters()
    dev = "cuda"
    got = [model(prompt.view(1, -1).to(dev), torch.arange(T, device=dev))[0, -1].float().cpu()]
    for i in range(n_total - T - 1):
        got.append(model(toks32[T + i].view(1, 1).to(dev), torch.tensor([T + i], device=dev))[0, -1].float().cpu())
    got = torch.stack(got)
    scale = logits32.abs().max().item()
    e_ref = (logits16 - logits32).abs().max().item()
    e_k = (got - logits32).abs().max().item()
    tol = 2.0 * e_ref + 1e-3 * scale
    print(f"[{label}] max|logit|={scale:.3f} bf16-reference err={e_ref:.4f} cuda err={e_k:.4f} tol={tol:.4f}")
    assert e_k <= tol
    top2 = logits32.topk(2, dim=-1).values
    decisive = (top2[:, 0] - top2[:, 1]) > tol
    agree = got.argmax(-1) == logits32.argmax(-1)
    assert bool(agree[decisive].all()), f"top-1 mismatch on decisive steps: {agree.tolist()}"
    print(f"[{label}] top-1 agreement {int(agree.sum())}/{len(agree)} (decisive steps {int(decisive.sum())})")
    # free-running greedy generate() through the CUDA graph == the eager module loop of the same model
    for blk in model.transformer.h:
        blk.attn.kv_cache.reset_parameters()
    out = gb.generate(model, prompt.to(dev), n_total, temperature=0.0)
    assert out.shape[0] == n_total and torch.equal(out[:T].cpu(), prompt)
    first_div = next((i for i in range(T, n_total) if out[i].item() != toks32[i].item()), None)
    if first_div is not None:  # a divergence is only legitimate at a non-decisive step
        assert not bool(decisive[first_div - T]), f"free-running rollout diverged at decisive step {first_div}"
    return out


@pytest.mark.parametrize("name", ["llama_gqa_hs64", "llama_mha_hs128"])
@pytest.mark.parametrize("mode", ["nf4", "fp4", "int4"])
@torch.inference_mode()
def test_quantized_toy_against_oracle(golden, name, mode):
    g = golden(name)
    cfgd, kw = cfg_from_golden(g)
    rc = M.RefConfig.from_any(cfgd)
    p = M.init_weights(rc, seed=1234)
    model = build_cuda_model(kw, p, mode, max_seq=40)
    assert model.fused_decode_ready()
    pw = oracle_weights(p, mode)
    # dequantised weights are bit-exact to the oracle's
    for k in _linear_keys(p):
        mod = model.get_submodule(k[: -len(".weight")])
        assert torch.equal(mod.dequantize(torch.float32).cpu(), pw[k]), k
    prompt = torch.from_numpy(g["gen_prompt"])
    check_rollout(model, rc, pw, prompt, 30, f"{name}/{mode}")


@torch.inference_mode()
def test_llama7b_width_two_layers_nf4():
    """Llama-2-7B shapes (d=4096, 32 heads, I=11008, V=32000) at depth 2: every kernel at its BASELINE tile shapes."""
    from lit_gpt_b200 import Config

    cfg = Config.from_name("Llama-2-7b-hf", n_layer=2)
    rc = M.RefConfig.from_any(cfg)
    p = M.init_weights(rc, seed=1234, dtype=torch.bfloat16)
    kw = dict(name="Llama-2-7b-hf")
    from lit_gpt_b200.config import name_to_config

    kw = dict(name_to_config["Llama-2-7b-hf"], n_layer=2)
    model = build_cuda_model(kw, p, "nf4", max_seq=48)
    assert model.fused_decode_ready()
    # every Linear of the headline shape runs on the streaming kernel (a silent fall-back to the generic GEMV is a perf bug)
    blk = model.transformer.h[0]
    assert all(l.streams(1) for l in (blk.attn.attn, blk.attn.proj, blk.mlp.proj, model.lm_head)) and blk.mlp.fc_1.streams(1, swiglu=True)
    pw = oracle_weights(p, "nf4")
    g = torch.Generator().manual_seed(1234)
    prompt = torch.randint(0, cfg.vocab_size, (16,), generator=g)
    check_rollout(model, rc, pw, prompt, 40, "llama7b-width-2L/nf4")


@torch.inference_mode()
def test_mistral7b_width_two_layers_int4_gqa():
    """Mistral-7B shapes (d = 4096, 32 heads / 8 KV groups, I = 14336, BASELINE config 4) at depth 2, gptq.int4: the grouped-
    query decode kernel, the int4 streaming GEMV at K = 14336 and the prefill GEMM, end to end against the oracle."""
    from lit_gpt_b200 import Config
    from lit_gpt_b200.config import name_to_config

    cfg = Config.from_name("Mistral-7B-v0.1", n_layer=2)
    rc = M.RefConfig.from_any(cfg)
    p = M.init_weights(rc, seed=77, dtype=torch.bfloat16)
    kw = dict(name_to_config["Mistral-7B-v0.1"], n_layer=2)
    model = build_cuda_model(kw, p, "int4", max_seq=48)
    assert model.fused_decode_ready() and model.config.n_query_groups == 8
    blk = model.transformer.h[0]
    assert all(l.streams(1) for l in (blk.attn.attn, blk.attn.proj, blk.mlp.proj, model.lm_head)) and blk.mlp.fc_1.streams(1, swiglu=True)
    pw = oracle_weights(p, "int4")
    g = torch.Generator().manual_seed(4)
    prompt = torch.randint(0, cfg.vocab_size, (12,), generator=g)
    check_rollout(model, rc, pw, prompt, 36, "mistral7b-width-2L/int4")


@pytest.mark.parametrize("B", [4, 16])
@torch.inference_mode()
def test_llama13b_width_batched_decode_int4(B):
    """Llama-2-13B shapes (d = 5120, I = 13824, BASELINE config 3) at depth 2, gptq.int4, batch 4 and 16: every row of the
    batched decode step (streaming GEMV: k-windows at batch 16) against the oracle run row by row."""
    from lit_gpt_b200 import Config
    from lit_gpt_b200.config import name_to_config

    cfg = Config.from_name("Llama-2-13b-hf", n_layer=2)
    rc = M.RefConfig.from_any(cfg)
    p = M.init_weights(rc, seed=13, dtype=torch.bfloat16)
    kw = dict(name_to_config["Llama-2-13b-hf"], n_layer=2)
    model = build_cuda_model(kw, p, "int4", max_seq=16, batch=B)
    assert model.fused_decode_ready()
    blk = model.transformer.h[0]
    assert all(l.streams(B) for l in (blk.attn.attn, blk.attn.proj, blk.mlp.proj, model.lm_head)) and blk.mlp.fc_1.streams(B, swiglu=True)
    pw = oracle_weights(p, "int4")
    g = torch.Generator().manual_seed(5)
    T = 6
    prompts = torch.randint(0, cfg.vocab_size, (B, T), generator=g)
    nxt = torch.randint(0, cfg.vocab_size, (B, 1), generator=g)
    model(prompts.cuda(), torch.arange(T, device="cuda"))
    got = model(nxt.cuda(), torch.tensor([T], device="cuda"))[:, -1].float().cpu()   # one fused decode step, B rows
    cos, sin = model.cos.cpu(), model.sin.cpu()
    rows = [0, B // 2, B - 1] if B > 4 else list(range(B))
    for r in rows:
        outs = {}
        for dtype, c, s_ in ((torch.float32, cos.float(), sin.float()), (torch.bfloat16, cos, sin)):
            st = M.RefState(rc, 16, 1, dtype, cos=c, sin=s_)
            pd = {k: v.to(dtype) for k, v in pw.items()}
            M.gpt_forward(rc, pd, prompts[r : r + 1], torch.arange(T), st)
            outs[dtype] = M.gpt_forward(rc, pd, nxt[r : r + 1], torch.tensor([T]), st)[0, -1].float()
        scale = outs[torch.float32].abs().max().item()
        e_ref = (outs[torch.bfloat16] - outs[torch.float32]).abs().max().item()
        e_k = (got[r] - outs[torch.float32]).abs().max().item()
        tol = 2.0 * e_ref + 1e-3 * scale
        print(f"[llama13b-width-2L/int4/B={B}/row {r}] bf16-reference err={e_ref:.4f} cuda err={e_k:.4f} tol={tol:.4f}")
        assert e_k <= tol, (r, e_k, tol)


@torch.inference_mode()
def test_long_context_rollout_bf16_rope_tables(golden):
    """A rollout past position 256 with the reference's bf16-default RoPE tables (build_rope_cache runs under the
    bf16 default dtype in generate/base.py, so the tables themselves are bf16): the decode attention's split / cluster paths at
    a context of ~300 and the table handling, against the oracle using the same tables."""
    g = golden("llama_mha_hs128")
    cfgd, kw = cfg_from_golden(g)
    kw = dict(kw, block_size=512)
    cfgd = dict(cfgd, block_size=512)
    rc = M.RefConfig.from_any(cfgd)
    p = M.init_weights(rc, seed=99)
    model = build_cuda_model(kw, p, "nf4", max_seq=330)
    # generate/base.py builds the tables under the bf16 default dtype: bf16 tables (positions past 256 are not bf16-exact)
    model.cos, model.sin = model.cos.to(torch.bfloat16), model.sin.to(torch.bfloat16)
    model.__dict__.pop("_decode_graphs", None)
    pw = oracle_weights(p, "nf4")
    gen = torch.Generator().manual_seed(6)
    prompt = torch.randint(0, cfgd["vocab_size"], (280,), generator=gen)
    check_rollout(model, rc, pw, prompt, 310, "long-context/nf4")


@torch.inference_mode()
def test_chat_stream_matches_generate(golden):
    """chat/base.py:24-68 on the fused path: the streamed tokens are generate()'s tokens, and a stop sequence ends the stream
    without being emitted."""
    import lit_gpt_b200.chat.base as chat
    import lit_gpt_b200.generate.base as gb

    g = golden("llama_gqa_hs64")
    cfgd, kw = cfg_from_golden(g)
    p = M.init_weights(M.RefConfig.from_any(cfgd), seed=1234)
    model = build_cuda_model(kw, p, "nf4", max_seq=40)
    prompt = torch.from_numpy(g["gen_prompt"]).cuda()
    T = prompt.numel()
    full = gb.generate(model, prompt, 30, temperature=0.0)
    for blk in model.transformer.h:
        blk.attn.kv_cache.reset_parameters()
    stream = torch.cat(list(chat.generate(model, prompt, 30, temperature=0.0)))
    assert torch.equal(stream, full[T:])
    stop = [int(full[T + 9])]   # (multi-token hold-back rules: the reference's own vectors in tests/test_chat_cpu.py)
    first = next(i for i in range(T, 30) if int(full[i]) == stop[0])
    for blk in model.transformer.h:
        blk.attn.kv_cache.reset_parameters()
    cut = list(chat.generate(model, prompt, 30, temperature=0.0, stop_tokens=(stop,)))
    assert (torch.cat(cut).tolist() if cut else []) == full[T:first].tolist()


@torch.inference_mode()
def test_batched_decode_matches_single(golden):
    """batch 1/4/16 at the model level (generate() is batch-1): rows are independent."""
    g = golden("llama_mha_hs128")
    cfgd, kw = cfg_from_golden(g)
    rc = M.RefConfig.from_any(cfgd)
    p = M.init_weights(rc, seed=1234)
    outs = {}
    for B in (1, 4, 16):
        model = build_cuda_model(kw, p, "nf4", max_seq=32, batch=B)
        x = torch.arange(5).repeat(B, 1).cuda() + torch.arange(B).cuda()[:, None]
        model(x, torch.arange(5, device="cuda"))
        y = model(torch.full((B, 1), 7, device="cuda"), torch.tensor([5], device="cuda"))
        outs[B] = y.float().cpu()
    assert (outs[4][0] - outs[1][0]).abs().max() <= 2 ** -6 * outs[1].abs().max()
    assert (outs[16][0] - outs[1][0]).abs().max() <= 2 ** -6 * outs[1].abs().max()
    assert (outs[16][:4] - outs[4]).abs().max() <= 2 ** -6 * outs[4].abs().max()


@torch.inference_mode()
def test_sampling_path_in_graph(golden):
    import lit_gpt_b200.generate.base as gb

    g = golden("llama_gqa_hs64")
    cfgd, kw = cfg_from_golden(g)
    p = M.init_weights(M.RefConfig.from_any(cfgd), seed=1234)
    model = build_cuda_model(kw, p, "nf4", max_seq=40)
    prompt = torch.from_numpy(g["gen_prompt"]).cuda()
    from lit_gpt_b200.fabric import seed_everything

    seed_everything(1234)
    out = gb.generate(model, prompt, 30, temperature=0.8, top_k=20)
    assert out.shape[0] == 30 and int(out.max()) < cfgd["padded_vocab_size"]
    # a second sample continues the device's ONE Philox stream (prefill draw and captured graph share it), a re-seed replays it
    out_b = gb.generate(model, prompt, 30, temperature=0.8, top_k=20)
    seed_everything(1234)
    out_c = gb.generate(model, prompt, 30, temperature=0.8, top_k=20)
    assert torch.equal(out, out_c) and not torch.equal(out, out_b)
    # eos stops generation (polled) and the eos token itself is kept, as in the reference loop
    greedy = gb.generate(model, prompt, 30, temperature=0.0)
    eos = int(greedy[12])
    cut = gb.generate(model, prompt, 30, temperature=0.0, eos_id=eos)
    first = next(i for i in range(6, 30) if int(greedy[i]) == eos)
    assert cut.tolist() == greedy[: first + 1].tolist()
    # short prompts run their prefill as a captured graph (generate/base.py PrefillGraph): same tokens as the eager prefill, greedy
    # and sampled (the graph's warm-up draws must not advance the Philox stream), also when replayed with another prompt
    assert prompt.numel() <= gb.PREFILL_GRAPH_MAX_T and model.__dict__.get("_prefill_graphs")
    prompt2 = (prompt + 3) % cfgd["vocab_size"]
    seed_everything(77)
    with_graph = [gb.generate(model, q, 30, temperature=t, top_k=k) for q, t, k in ((prompt, 0.0, None), (prompt2, 0.0, None), (prompt2, 0.8, 20), (prompt, 0.8, 20))]
    keep = gb.PREFILL_GRAPH_MAX_T
    try:
        gb.PREFILL_GRAPH_MAX_T = 0
        seed_everything(77)
        eager = [gb.generate(model, q, 30, temperature=t, top_k=k) for q, t, k in ((prompt, 0.0, None), (prompt2, 0.0, None), (prompt2, 0.8, 20), (prompt, 0.8, 20))]
    finally:
        gb.PREFILL_GRAPH_MAX_T = keep
    for a, b in zip(with_graph, eager):
        assert torch.equal(a, b)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs (the one-GPU variant is tests/test_gpu_tp_shared.py)")
@pytest.mark.parametrize("mode,inter", [("nf4", 0), ("int4", 0), ("nf4", 2112)])
def test_tp2_matches_oracle_multi_gpu(mode, inter):
    """Two ranks on two GPUs over NCCL / NVLink: the sharded model (peer-memory hand-off in the GEMVs) against the oracle on
    identically sharded-then-quantized weights, identical logits on every rank, and agreement with the all-NCCL path
    (B200_TP_NCCL=1).  inter = 2112: the mlp.proj shard (1056 columns) is not block-aligned -> folded tiled layout."""
    import os
    import subprocess
    import sys

    root = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
    env = dict(os.environ, B200_TP_SHARED_DEVICE="0", B200_TP_NCCL="0", OMP_NUM_THREADS="4")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29631", os.path.join(root, "tests", "tp_shared_worker.py"), root, mode, "8", "4", "0", str(inter)]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0 and "TP_SHARED_OK" in out.stdout, out.stdout[-2000:] + out.stderr[-4000:]
    print(out.stdout[-500:])


@pytest.mark.parametrize("mode", ["nf4", "int4", "nf4-dq"])
@torch.inference_mode()
def test_single_resident_copy(golden, mode):
    """`GPT.release_reference_layouts()`: only the tiled copy of every 4-bit Linear stays in HBM (half the bytes); prefill, decode,
    generate(), dequantize() and state_dict() give exactly what they gave before (the reference layout is rebuilt on the fly by the
    exact inverse of the tiling)."""
    import lit_gpt_b200.generate.base as gb
    from lit_gpt_b200 import QLinear

    g = golden("llama_mha_hs128")
    cfgd, kw = cfg_from_golden(g)
    p = M.init_weights(M.RefConfig.from_any(cfgd), seed=21)
    model = build_cuda_model(kw, p, mode, max_seq=40)
    prompt = torch.from_numpy(g["gen_prompt"]).cuda()
    T = prompt.numel()

    def run():
        for blk in model.transformer.h:
            blk.attn.kv_cache.reset_parameters()
        y = model(prompt.view(1, -1), torch.arange(T, device="cuda")).clone()
        d = model(torch.tensor([[7]], device="cuda"), torch.tensor([T], device="cuda")).clone()
        for blk in model.transformer.h:
            blk.attn.kv_cache.reset_parameters()
        return y, d, gb.generate(model, prompt, 30, temperature=0.0).clone()

    def weight_bytes():
        n = 0
        for m in model.modules():
            if isinstance(m, QLinear):
                for t in (m.weight.data, m.absmax, m.sz, m.tiled):
                    n += 0 if t is None else t.numel() * t.element_size()
        return n

    before = run()
    sd0 = {k: v.clone() for k, v in model.state_dict().items()}
    dq0 = model.transformer.h[0].mlp.fc_1.dequantize(torch.float32).clone()
    b0 = weight_bytes()
    n = model.release_reference_layouts()
    assert n == sum(isinstance(m, QLinear) for m in model.modules())
    assert weight_bytes() <= 0.52 * b0
    after = run()
    assert all(torch.equal(a, b) for a, b in zip(before, after))
    sd1 = model.state_dict()
    assert sd0.keys() == sd1.keys() and all(torch.equal(sd0[k].view(torch.uint8).flatten(), sd1[k].view(torch.uint8).flatten()) for k in sd0)
    assert torch.equal(dq0, model.transformer.h[0].mlp.fc_1.dequantize(torch.float32))


def _sequential_model(kw, p, mode, devices, max_seq):
    """The load flow of generate/sequentially.py:171-190: weights on the CPU, QLinear modules swapped in, then placed."""
    from lit_gpt_b200 import GPT, Config, QuantPrecision
    from lit_gpt_b200.generate.sequentially import sequential

    prev = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    try:
        with torch.device("meta"):
            model = GPT(Config(**kw))
    finally:
        torch.set_default_dtype(prev)
    model.load_state_dict({k: v.to(torch.bfloat16) for k, v in p.items()}, assign=True)
    model = QuantPrecision(mode, torch.bfloat16).convert_module(model).eval()
    return sequential(model, torch.device("cuda", 0), max_seq, devices)


@pytest.mark.parametrize("devices", [1, 2])
@torch.inference_mode()
def test_sequential_placement_matches_single_device(golden, devices):
    """generate/sequentially.py:31-97: blocks placed over `devices` GPUs by one process give the logits and the
    greedy tokens of the same quantized model on one GPU, bit for bit (same kernels, the residual stream hops devices)."""
    if torch.cuda.device_count() < devices:
        pytest.skip(f"needs {devices} GPUs")
    import lit_gpt_b200.generate.base as gb
    from lit_gpt_b200.generate.sequentially import move_block_input, move_block_output

    g = golden("llama_gqa_hs64")
    cfgd, kw = cfg_from_golden(g)
    p = M.init_weights(M.RefConfig.from_any(cfgd), seed=1234)
    single = build_cuda_model(kw, p, "nf4", max_seq=40)
    prompt = torch.from_numpy(g["gen_prompt"]).cuda()
    T = prompt.numel()
    y1 = single(prompt.view(1, -1), torch.arange(T, device="cuda")).float().cpu()
    d1 = single(torch.tensor([[7]], device="cuda"), torch.tensor([T], device="cuda")).float().cpu()
    for blk in single.transformer.h:
        blk.attn.kv_cache.reset_parameters()
    t1 = gb.generate(single, prompt, 30, temperature=0.0).cpu()
    model = _sequential_model(kw, p, "nf4", devices, 40)
    n_layer = cfgd["n_layer"]
    devs = [str(b.attn.attn.weight.device) for b in model.transformer.h]
    assert devs == [f"cuda:{i // (n_layer // devices)}" for i in range(n_layer)]
    assert all(str(b.attn.kv_cache.k.device) == d for b, d in zip(model.transformer.h, devs))
    assert str(model.lm_head.weight.device) == "cuda:0" and model.max_seq_length == 40
    if devices > 1:   # the reference's hook placement (tests/test_generate_sequentially.py:157-260)
        last = model.transformer.h[-1]
        assert [h.func for h in last._forward_hooks.values()] == [move_block_output]
        assert all(([h.func for h in b._forward_pre_hooks.values()] == [move_block_input]) == (i >= n_layer // devices)
                   for i, b in enumerate(model.transformer.h))
        assert model.spans_devices()
    assert model.fused_decode_ready()
    y2 = model(prompt.view(1, -1), torch.arange(T, device="cuda")).float().cpu()
    d2 = model(torch.tensor([[7]], device="cuda"), torch.tensor([T], device="cuda")).float().cpu()
    assert torch.equal(y1, y2) and torch.equal(d1, d2)
    for blk in model.transformer.h:
        blk.attn.kv_cache.reset_parameters()
    t2 = gb.generate(model, prompt, 30, temperature=0.0).cpu()
    assert torch.equal(t1, t2)


@torch.inference_mode(False)
def test_traced_forward_has_no_graph_breaks():
    """Mirror of the reference's tests/test_model.py:564-575: dynamo traces GPT.forward (prefill and the cached decode call)
    without graph breaks, because every kernel is a registered b200gpt::* custom op with a fake kernel (lit_gpt_b200/ops.py)."""
    import torch._dynamo

    from lit_gpt_b200 import Config

    kw = dict(name="trace", block_size=64, vocab_size=512, padded_vocab_size=512, n_layer=2, n_head=4, n_query_groups=2, n_embd=256,
              rotary_percentage=1.0, parallel_residual=False, bias=False, _norm_class="RMSNorm", _mlp_class="LLaMAMLP", intermediate_size=512)
    p = M.init_weights(M.RefConfig.from_any(Config(**kw)), seed=3, dtype=torch.bfloat16)
    model = build_cuda_model(kw, p, "nf4", max_seq=32)
    x = torch.tensor([[3, 1, 4, 1, 5]], device="cuda")
    with torch.no_grad():
        for inp, pos in ((x, torch.arange(5, device="cuda")), (x[:, :1], torch.tensor([5], device="cuda"))):
            torch._dynamo.reset()
            ex = torch._dynamo.explain(model)(inp, pos)
            assert ex.graph_break_count == 0 and ex.graph_count == 1, (ex.break_reasons, ex.graph_count)
