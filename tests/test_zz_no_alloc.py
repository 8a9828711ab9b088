"""b200_attn_decode_noalloc (include/b200gpt.h): the allocation-free variant of the decode attention, whose split-K scratch
lives in a static device array inside the library.  It has no counterpart in the reference,
so the checker here is the existing entry point b200_attn_decode on the same inputs.

Kept in a file of its own that pytest collects last."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _call(lib, C, fn, qkv, cos, sin, pos, kc, vc, H, G, hs, splits, out, ws=None):
    B, max_seq = kc.shape[0], kc.shape[2]
    args = C.AttnDecodeArgs(
        qkv=qkv.data_ptr(), cos=cos.data_ptr(), sin=sin.data_ptr(), table_dtype=C.F32, table_preselected=0, rope_n_elem=hs,
        pos=pos.data_ptr(), pos_is_int64=1, k_cache=kc.data_ptr(), v_cache=vc.data_ptr(), out=out.data_ptr(), batch=B, n_head=H,
        n_groups=G, head_size=hs, max_seq=max_seq, num_splits=splits, workspace=None if ws is None else ws.data_ptr(),
        workspace_bytes=0 if ws is None else ws.numel())
    rc = getattr(lib, fn)(ctypes.byref(args), C.stream_ptr(qkv.device))
    return rc, out


@pytest.mark.parametrize("H,G,splits", [(32, 32, 16), (32, 8, 0), (8, 1, 5)])  # workspace-merged scalar kernel, grouped-query kernel (auto), MQA
def test_attn_decode_noalloc_matches_attn_decode(H, G, splits):
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from lit_gpt_b200 import _cabi as C
    from oracle import model_ref as M

    lib = C.load()
    hs, B, max_seq = 128, 2, 1024
    assert lib.b200_attn_decode_noalloc_max_bytes() == 4 << 20
    g = torch.Generator().manual_seed(H + G)
    cos, sin = M.build_rope_cache(max_seq, hs, pos_dtype=torch.float32)
    cos, sin = cos.float().cuda().contiguous(), sin.float().cuda().contiguous()
    ws = torch.zeros(lib.b200_attn_decode_workspace_bytes(B, H, hs, 64), dtype=torch.uint8, device="cuda")
    for it, p in enumerate((700, 701)):  # two calls in a row share the built-in scratch
        qkv = torch.randn(B, (H + 2 * G) * hs, generator=g).bfloat16().cuda()
        kc = torch.randn(B, G, max_seq, hs, generator=g).bfloat16()
        vc = torch.randn(B, G, max_seq, hs, generator=g).bfloat16()
        kc[:, :, p:], vc[:, :, p:] = 0, 0
        pos = torch.tensor([p], device="cuda")
        k1, v1, k2, v2 = kc.cuda(), vc.cuda(), kc.cuda(), vc.cuda()
        want, got = (torch.empty((B, H * hs), dtype=torch.bfloat16, device="cuda") for _ in range(2))  # caller-owned outputs
        rc, _ = _call(lib, C, "b200_attn_decode", qkv, cos, sin, pos, k1, v1, H, G, hs, splits, want, ws)
        assert rc == 0, lib.b200_last_error()
        torch.cuda.synchronize()
        before = (torch.cuda.memory_allocated(), torch.cuda.memory_reserved(), torch.cuda.mem_get_info()[0])
        rc, _ = _call(lib, C, "b200_attn_decode_noalloc", qkv, cos, sin, pos, k2, v2, H, G, hs, splits, got)
        assert rc == 0, lib.b200_last_error()
        torch.cuda.synchronize()
        # nothing was allocated: neither through torch's allocator nor behind its back (free device memory unchanged)
        assert (torch.cuda.memory_allocated(), torch.cuda.memory_reserved()) == before[:2]
        if it > 0:  # (the first call of a process may make the CUDA runtime load the kernel's module lazily)
            assert torch.cuda.mem_get_info()[0] == before[2]
        assert torch.equal(k1, k2) and torch.equal(v1, v2), "cache append differs"
        tol = 2 ** -7 * max(1.0, want.float().abs().max().item())  # the bar of tests/test_gpu_kernels.py::test_attn_decode
        assert (got.float() - want.float()).abs().max().item() <= tol, f"call {it}"


def test_attn_decode_noalloc_rejects_oversized_requests():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from lit_gpt_b200 import _cabi as C

    lib = C.load()
    # 16 sequences x 64 heads x 64 splits needs ~33 MB of scratch: beyond the 4 MiB array -> a clear error, nothing launched
    args = C.AttnDecodeArgs(batch=16, n_head=64, n_groups=8, head_size=128, max_seq=4096, num_splits=0)
    rc = lib.b200_attn_decode_noalloc(ctypes.byref(args), None)
    assert rc == 3  # B200_ERR_UNSUPPORTED
    assert b"scratch" in lib.b200_last_error() and b"caller-owned workspace" in lib.b200_last_error()
