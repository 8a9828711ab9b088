"""Fused on-device sampler (csrc/sample.cu) against the reference formulation of generate/base.py:22-41."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def K():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from lit_gpt_b200 import kernels

    return kernels


def _ref_sample(logits, temperature, top_k, noise):
    """sample() of generate/base.py:29-41 on row [0, -1] in fp32, with the Exp(1) draw of base.py:23-26 on given noise."""
    l = logits[0, -1].float()
    if top_k is not None:
        v, i = torch.topk(l, min(top_k, l.size(-1)))
        l = torch.full_like(l, float("-inf")).scatter_(-1, i, v)
    if temperature > 0.0:
        probs = torch.softmax(l / temperature, dim=-1)
        return int(torch.argmax(probs / noise, dim=-1))
    return int(torch.argmax(l, dim=-1))


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("V", [32000, 50304, 1000, 37])
@pytest.mark.parametrize("temperature,top_k", [(0.8, 200), (1.0, None), (0.7, 1), (2.0, 5), (0.0, None), (0.0, 50), (1.3, 100000)])
def test_sample_given_noise_matches_reference(K, dtype, V, temperature, top_k):
    g = torch.Generator().manual_seed(V + (top_k or 0))
    for trial in range(4):
        logits = (torch.randn(1, 3, V, generator=g) * 3).to(dtype)
        if dtype == torch.float32:  # distinct values: the kept set is unambiguous (bf16 ties are covered below)
            logits += torch.arange(V).view(1, 1, V) * 1e-6
        noise = torch.empty(V).exponential_(1, generator=g)
        want = _ref_sample(logits, temperature, top_k, noise)
        tok = torch.zeros(1, dtype=torch.int64, device="cuda")
        pos = torch.tensor([7], device="cuda")
        hist = torch.zeros(16, dtype=torch.int64, device="cuda")
        K.sample_(logits.cuda(), tok, temperature, top_k, None, pos_inout=pos, history=hist, noise=noise.cuda())
        if dtype == torch.bfloat16 and top_k is not None and 1 < top_k < V and temperature > 0:
            # ties at the threshold: torch.topk's choice among equal values is unspecified; the token must be one of the
            # top_k largest VALUES and, when it is strictly above the threshold, identical
            l = logits[0, -1].float()
            thr = torch.topk(l, top_k).values[-1]
            assert l[tok.item()] >= thr
            if l[want] > thr and l[tok.item()] > thr and (l == thr).sum() == 1:
                assert tok.item() == want
        else:
            assert tok.item() == want, (trial, tok.item(), want)
        assert pos.item() == 8 and hist[8].item() == tok.item() and hist.sum().item() == tok.item()


def test_sample_tie_rule_and_counts(K):
    """bf16-style ties: exactly top_k entries are kept, ties at the threshold in index order."""
    V, k = 64, 5
    logits = torch.zeros(1, 1, V)
    logits[0, 0, [3, 9]] = 2.0          # two above the threshold
    logits[0, 0, [10, 20, 30, 40, 50]] = 1.0   # five tie at the threshold, only the first three (10, 20, 30) are kept
    seen = set()
    for i in range(V):
        noise = torch.full((V,), 1e6)
        noise[i] = 1e-6     # forces index i if it is in the kept set
        tok = torch.zeros(1, dtype=torch.int64, device="cuda")
        K.sample_(logits.cuda(), tok, 1.0, k, None, noise=noise.cuda())
        if tok.item() == i:
            seen.add(i)
    assert seen == {3, 9, 10, 20, 30}


def test_sample_philox_distribution_and_replay(K):
    """Philox draws: empirical frequencies match softmax(top_k(l) / T); a captured graph draws fresh numbers per replay;
    equal (seed, offset) reproduce the same token."""
    V, T, k = 1000, 0.9, 8
    g = torch.Generator().manual_seed(0)
    logits = (torch.randn(1, 1, V, generator=g) * 2).cuda()
    l = logits[0, 0].cpu()
    v, i = torch.topk(l, k)
    p = torch.zeros(V)
    p[i] = torch.softmax(v / T, -1)
    rng = K.new_rng_state(torch.device("cuda"), seed=1234)
    tok = torch.zeros(1, dtype=torch.int64, device="cuda")
    hist = torch.zeros(20001, dtype=torch.int64, device="cuda")
    pos = torch.tensor([0], device="cuda")
    K.sample_(logits, tok, T, k, rng, pos_inout=pos, history=hist)   # warm-up outside capture
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        K.sample_(logits, tok, T, k, rng, pos_inout=pos, history=hist)
    n = 19999
    for _ in range(n):
        graph.replay()
    torch.cuda.synchronize()
    assert rng[1].item() == n + 1 and pos.item() == n + 1
    draws = hist[1 : n + 2].cpu()
    freq = torch.bincount(draws, minlength=V).float() / draws.numel()
    assert freq[p == 0].sum() == 0, "a token outside the top-k was drawn"
    assert (freq - p).abs().max() < 0.015, (freq[i], p[i])
    # reproducibility: same (seed, offset) -> same token
    rng2 = torch.tensor([1234, 5], dtype=torch.int64, device="cuda")
    K.sample_(logits, tok, T, k, rng2)
    assert tok.item() == draws[5].item()


def test_generate_sampling_stays_on_device(K):
    """generate() with temperature / top_k: graph-captured fused sampler; seeds reproduce, tokens stay inside the top-k."""
    import lit_gpt_b200.generate.base as gb
    from test_gpu_model import build_cuda_model
    from lit_gpt_b200 import Config
    from oracle import model_ref as M

    kw = dict(name="s", block_size=64, vocab_size=512, padded_vocab_size=512, n_layer=2, n_head=4, n_query_groups=2, n_embd=256,
              rotary_percentage=1.0, parallel_residual=False, bias=False, _norm_class="RMSNorm", _mlp_class="LLaMAMLP", intermediate_size=512)
    p = M.init_weights(M.RefConfig.from_any(Config(**kw)), seed=7, dtype=torch.bfloat16)
    model = build_cuda_model(kw, p, "nf4", max_seq=40)
    prompt = torch.tensor([1, 2, 3, 4, 5], device="cuda")
    outs = []
    for _ in range(2):
        torch.manual_seed(99)
        for blk in model.transformer.h:
            blk.attn.kv_cache.reset_parameters()
        model.__dict__.pop("_decode_graphs", None)
        gb._RNG.clear()
        outs.append(gb.generate(model, prompt, 30, temperature=0.8, top_k=20).cpu())
    assert torch.equal(outs[0], outs[1]) and outs[0].numel() == 30
    torch.manual_seed(100)
    for blk in model.transformer.h:
        blk.attn.kv_cache.reset_parameters()
    model.__dict__.pop("_decode_graphs", None)
    other = gb.generate(model, prompt, 30, temperature=0.8, top_k=20).cpu()
    assert not torch.equal(other, outs[0])
