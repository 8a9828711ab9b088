"""Tensor-level wrappers over the C ABI (``include/b200gpt.h``): argument checking, output allocation, workspaces.

Every function launches on torch's current CUDA stream and never synchronises, so calls are capturable in a CUDA
graph.  CUDA tensors only — no CPU path exists (``B200Error`` is raised instead).
"""
from __future__ import annotations

import ctypes
import functools
from typing import Dict, Optional, Tuple

import torch
import torch._dynamo

from . import _cabi as C

QTYPES = {"nf4": C.Q_NF4, "fp4": C.Q_FP4, "int4": C.Q_INT4}
_workspaces: Dict[Tuple[str, int], torch.Tensor] = {}


def _ws(kind: str, device: torch.device, nbytes: int) -> torch.Tensor:
    """Persistent zero-initialised workspace per (kind, device).  Split-K counters inside are self-cleaning."""
    key = (kind, device.index if device.index is not None else torch.cuda.current_device())
    buf = _workspaces.get(key)
    if buf is None or buf.numel() < nbytes:
        if torch.cuda.is_current_stream_capturing():
            raise C.B200Error(f"workspace '{kind}' must be allocated before CUDA-graph capture (run one eager step first)")
        buf = torch.zeros(nbytes, dtype=torch.uint8, device=device)
        _workspaces[key] = buf
    return buf


def _bf16c(t: Optional[torch.Tensor], name: str) -> Optional[torch.Tensor]:
    if t is None:
        return None
    if t.dtype != torch.bfloat16:
        raise C.B200Error(f"{name} must be bfloat16 (got {t.dtype}); the kernels compute in bf16 with fp32 accumulation")
    return t if t.is_contiguous() else t.contiguous()


# ---------------------------------------------------------------------------------------------------------
# quantisation
# ---------------------------------------------------------------------------------------------------------
def quantize_4bit(w: torch.Tensor, quant_type: str = "nf4", blocksize: int = 64) -> Tuple[torch.Tensor, torch.Tensor]:
    """-> (packed uint8 [(numel+1)//2, 1], absmax fp32 [ceil(numel/blocksize)]) — bitsandbytes Params4bit layout."""
    C.require_cuda(w)
    if w.dtype not in C.DTYPE_CODE:
        raise C.B200Error(f"quantize_4bit: unsupported dtype {w.dtype}")
    w = w.contiguous()
    n = w.numel()
    packed = torch.empty(((n + 1) // 2, 1), dtype=torch.uint8, device=w.device)
    absmax = torch.empty(((n + blocksize - 1) // blocksize,), dtype=torch.float32, device=w.device)
    C.check(C.load().b200_quantize_4bit(w.data_ptr(), C.DTYPE_CODE[w.dtype], n, blocksize, QTYPES[quant_type],
                                        packed.data_ptr(), absmax.data_ptr(), C.stream_ptr(w.device)), "quantize_4bit")
    return packed, absmax


def dequantize_4bit(packed: torch.Tensor, absmax: torch.Tensor, shape, dtype: torch.dtype = torch.bfloat16,
                    quant_type: str = "nf4", blocksize: int = 64) -> torch.Tensor:
    C.require_cuda(packed, absmax)
    out = torch.empty(tuple(shape), dtype=dtype, device=packed.device)
    C.check(C.load().b200_dequantize_4bit(packed.data_ptr(), absmax.data_ptr(), out.numel(), blocksize, QTYPES[quant_type],
                                          out.data_ptr(), C.DTYPE_CODE[dtype], C.stream_ptr(packed.device)), "dequantize_4bit")
    return out


_dynamic_map_cache: Dict[int, torch.Tensor] = {}


def dynamic_map(device: torch.device) -> torch.Tensor:
    """bitsandbytes' 256-entry signed 8-bit "dynamic" code (functional.create_dynamic_map(signed=True, 7, 8)), sorted."""
    key = device.index if device.index is not None else torch.cuda.current_device()
    if key not in _dynamic_map_cache:
        import numpy as np

        f32 = np.float32
        data = []
        for i in range(7):
            items = int(2 ** i + 1)
            b = np.linspace(0.1, 1, items, dtype=f32)
            means = ((b[:-1] + b[1:]) / f32(2.0)).astype(f32)
            scale = 10 ** (-6 + i)
            data += (scale * means.astype(np.float64)).tolist() + (-scale * means.astype(np.float64)).tolist()
        data += [0, 1.0]
        data += [0] * (256 - len(data))
        data.sort()
        _dynamic_map_cache[key] = torch.tensor(np.array(data, dtype=f32), device=device)
    return _dynamic_map_cache[key]


def double_quant_absmax(absmax: torch.Tensor, offset: Optional[float] = None):
    """-> (q uint8 [n], absmax2 fp32 [ceil(n/256)], offset float, absmax_dequantised fp32 [n])  (SURVEY Q2)."""
    C.require_cuda(absmax)
    absmax = absmax.contiguous().float()
    n = absmax.numel()
    if offset is None:
        offset = float(absmax.mean())
    q = torch.empty(n, dtype=torch.uint8, device=absmax.device)
    a2 = torch.empty((n + 255) // 256, dtype=torch.float32, device=absmax.device)
    deq = torch.empty_like(absmax)
    code = dynamic_map(absmax.device)
    C.check(C.load().b200_double_quant_absmax(absmax.data_ptr(), n, code.data_ptr(), offset, q.data_ptr(), a2.data_ptr(), deq.data_ptr(),
                                              C.stream_ptr(absmax.device)), "double_quant_absmax")
    return q, a2, offset, deq


def dequant_absmax(qabs: torch.Tensor, absmax2: torch.Tensor, offset: float) -> torch.Tensor:
    """Load-time inverse of ``double_quant_absmax`` for pre-quantized checkpoints: fp32 absmax = code[q] * absmax2[i // 256]
    + offset with the kernel's two separate roundings (csrc/quantize.cu), so a reloaded module streams identical statistics."""
    C.require_cuda(qabs, absmax2)
    code = dynamic_map(qabs.device)
    n = qabs.numel()
    m = absmax2.float().repeat_interleave(256)[:n]
    return (code[qabs.long()] * m + torch.tensor(float(offset), dtype=torch.float32, device=qabs.device)).contiguous()


def int4_fuse_scale_zero(scales: torch.Tensor, qzeros: torch.Tensor) -> torch.Tensor:
    """Load-time rebuild of the kernel-side ``sz`` tensor, int32 (N, G) = half2{scale, zero} (low half: fp16 scale bits, high
    half: the 4-bit zero-point as fp16), from the checkpoint tensors of ``quantize_int4``.  Bit-exact to the quantizer's own."""
    N, G = scales.shape
    shifts = torch.arange(0, 32, 4, device=qzeros.device, dtype=torch.int32)
    z = ((qzeros.to(torch.int32).unsqueeze(-1) >> shifts) & 0xF).reshape(N, -1)[:, :G]
    lo = scales.contiguous().view(torch.int16).to(torch.int32) & 0xFFFF
    hi = z.to(torch.float16).view(torch.int16).to(torch.int32) & 0xFFFF
    return (lo | (hi << 16)).contiguous()


# ---------------------------------------------------------------------------------------------------------
# bnb.int8 (LLM.int8())
# ---------------------------------------------------------------------------------------------------------
def quantize_int8(w: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """-> (CB int8 (N, K), SCB fp32 (N)) — Linear8bitLt's row-wise weight quantisation (oracle/quant_ref.py:312-320)."""
    C.require_cuda(w)
    if w.dim() != 2 or w.dtype not in C.DTYPE_CODE:
        raise C.B200Error(f"quantize_int8 expects a 2-D floating-point weight (got {tuple(w.shape)}, {w.dtype})")
    w = w.contiguous()
    N, K = w.shape
    cb = torch.empty((N, K), dtype=torch.int8, device=w.device)
    scb = torch.empty((N,), dtype=torch.float32, device=w.device)
    C.check(C.load().b200_quantize_int8(w.data_ptr(), C.DTYPE_CODE[w.dtype], N, K, cb.data_ptr(), scb.data_ptr(),
                                        C.stream_ptr(w.device)), "quantize_int8")
    return cb, scb


def dequantize_int8(cb: torch.Tensor, scb: torch.Tensor, dtype: torch.dtype = torch.float16) -> torch.Tensor:
    C.require_cuda(cb, scb)
    N, K = cb.shape
    out = torch.empty((N, K), dtype=dtype, device=cb.device)
    C.check(C.load().b200_dequantize_int8(cb.data_ptr(), scb.data_ptr(), N, K, out.data_ptr(), C.DTYPE_CODE[dtype],
                                          C.stream_ptr(cb.device)), "dequantize_int8")
    return out


def linear_int8(x: torch.Tensor, cb: torch.Tensor, scb: torch.Tensor, bias: Optional[torch.Tensor] = None,
                threshold: float = 6.0) -> torch.Tensor:
    """LLM.int8() forward: x (..., K) fp16 -> (..., N) fp16 (MatMul8bitLt inference path, oracle/quant_ref.py:326-345)."""
    C.require_cuda(x, cb, scb, bias)
    if x.dtype != torch.float16:
        raise C.B200Error(f"linear_int8 computes in fp16 like bitsandbytes (got {x.dtype}); cast the activation first")
    N, K = cb.shape
    if x.shape[-1] != K:
        raise C.B200Error(f"linear_int8: activation has {x.shape[-1]} features, weight expects {K}")
    x2 = x.reshape(-1, K)
    x2 = x2 if x2.is_contiguous() else x2.contiguous()
    M = x2.shape[0]
    y = torch.empty((M, N), dtype=torch.float16, device=x.device)
    if M == 0:
        return y.reshape(*x.shape[:-1], N)
    if bias is not None and bias.dtype != torch.float16:
        bias = bias.to(torch.float16)
    lib = C.load()
    ws = _ws("int8", x.device, int(lib.b200_linear_int8_workspace_bytes(M, K)))
    a = C.LinearInt8Args(x=x2.data_ptr(), cb=cb.data_ptr(), scb=scb.data_ptr(), bias=None if bias is None else bias.data_ptr(),
                         y=y.data_ptr(), M=M, N=N, K=K, threshold=float(threshold), workspace=ws.data_ptr(),
                         workspace_bytes=ws.numel())
    C.check(lib.b200_linear_int8(ctypes.byref(a), C.stream_ptr(x.device)), "linear_int8")
    return y.reshape(*x.shape[:-1], N)


def quantize_int4(w: torch.Tensor, group: int = 128):
    """-> (qweight uint32-as-int32 (N, ceil(K/8)), scales fp16 (N, G), qzeros int32 (N, ceil(G/8)), sz int32 (N, G))."""
    C.require_cuda(w)
    if w.dim() != 2:
        raise C.B200Error("quantize_int4 expects a 2-D (out_features, in_features) weight")
    w = w.contiguous()
    N, K = w.shape
    G = (K + group - 1) // group
    qweight = torch.empty((N, (K + 7) // 8), dtype=torch.int32, device=w.device)
    scales = torch.empty((N, G), dtype=torch.float16, device=w.device)
    qzeros = torch.empty((N, (G + 7) // 8), dtype=torch.int32, device=w.device)
    sz = torch.empty((N, G), dtype=torch.int32, device=w.device)
    C.check(C.load().b200_quantize_int4(w.data_ptr(), C.DTYPE_CODE[w.dtype], N, K, group, qweight.data_ptr(), scales.data_ptr(),
                                        qzeros.data_ptr(), sz.data_ptr(), C.stream_ptr(w.device)), "quantize_int4")
    return qweight, scales, qzeros, sz


def dequantize_int4(qweight, scales, qzeros, K: int, group: int = 128, dtype: torch.dtype = torch.bfloat16) -> torch.Tensor:
    C.require_cuda(qweight, scales, qzeros)
    N = qweight.shape[0]
    out = torch.empty((N, K), dtype=dtype, device=qweight.device)
    C.check(C.load().b200_dequantize_int4(qweight.data_ptr(), scales.data_ptr(), qzeros.data_ptr(), N, K, group, out.data_ptr(),
                                          C.DTYPE_CODE[dtype], C.stream_ptr(qweight.device)), "dequantize_int4")
    return out


# ---------------------------------------------------------------------------------------------------------
# fused dequant-GEMV (batch <= 16)
# ---------------------------------------------------------------------------------------------------------
def qgemv(x: torch.Tensor, packed: torch.Tensor, scale: torch.Tensor, N: int, K: int, quant_type: str = "nf4",
          blocksize: int = 64, *, bias: Optional[torch.Tensor] = None, delta: Optional[torch.Tensor] = None,
          x_out: Optional[torch.Tensor] = None, norm_weight: Optional[torch.Tensor] = None, norm_eps: float = 1e-5,
          residual: Optional[torch.Tensor] = None, packed2: Optional[torch.Tensor] = None,
          scale2: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """y = epilogue(prologue(x) @ dequant(W).T); see b200_qgemv in include/b200gpt.h.  x: (..., K) with <= 16 rows."""
    C.require_cuda(x, packed, scale)
    x = _bf16c(x, "x")
    lead = x.shape[:-1]
    batch = x.numel() // K
    if x.shape[-1] != K:
        raise C.B200Error(f"qgemv: x has {x.shape[-1]} features, weight expects {K}")
    epilogue = C.EPI_SWIGLU if packed2 is not None else (C.EPI_RESIDUAL if residual is not None else C.EPI_STORE)
    y = out if out is not None else torch.empty((*lead, N), dtype=torch.bfloat16, device=x.device)
    nbytes = C.load().b200_qgemv_workspace_bytes()
    ws = _ws("qgemv", x.device, nbytes)
    delta, residual, bias, norm_weight = _bf16c(delta, "delta"), _bf16c(residual, "residual"), _bf16c(bias, "bias"), _bf16c(norm_weight, "norm_weight")
    args = C.QGemvArgs(
        packed=packed.data_ptr(), scale=scale.data_ptr(), packed2=C.ptr(packed2), scale2=C.ptr(scale2),
        qtype=QTYPES[quant_type], blocksize=blocksize, N=N, K=K, x=x.data_ptr(), delta=C.ptr(delta), x_out=C.ptr(x_out),
        norm_weight=C.ptr(norm_weight), norm_eps=norm_eps, batch=batch, epilogue=epilogue, bias=C.ptr(bias),
        residual=C.ptr(residual), y=y.data_ptr(), workspace=ws.data_ptr(), workspace_bytes=ws.numel())
    C.check(C.load().b200_qgemv(ctypes.byref(args), C.stream_ptr(x.device)), "qgemv")
    return y


# ---------------------------------------------------------------------------------------------------------
# tiled layout + streaming GEMV (the decode hot path)
# ---------------------------------------------------------------------------------------------------------
def tiled_bytes(N: int, K: int, quant_type: str = "nf4") -> int:
    """Size of the tiled buffer for an (N, K) matrix; 0 when the layout does not cover the shape."""
    return int(C.load().b200_tiled_bytes(N, K, QTYPES[quant_type]))


def tile_4bit(packed: torch.Tensor, scale: torch.Tensor, N: int, K: int, quant_type: str = "nf4", blocksize: int = 64,
              out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Reference-layout (packed, absmax) or (qweight, sz) -> flat uint8 tiled buffer (csrc/tile.cu)."""
    C.require_cuda(packed, scale)
    nbytes = tiled_bytes(N, K, quant_type)
    if nbytes == 0:
        raise C.B200Error(f"tile_4bit: shape ({N}, {K}) is not covered by the tiled layout for {quant_type}")
    if out is None:
        out = torch.empty(nbytes, dtype=torch.uint8, device=packed.device)
    C.check(C.load().b200_tile_4bit(packed.data_ptr(), scale.data_ptr(), N, K, QTYPES[quant_type], blocksize, out.data_ptr(),
                                    C.stream_ptr(packed.device)), "tile_4bit")
    return out


def untile_4bit(tiled: torch.Tensor, N: int, K: int, quant_type: str = "nf4", blocksize: int = 64):
    """Inverse of tile_4bit -> (packed, scale) in the reference layout, bit for bit."""
    C.require_cuda(tiled)
    if quant_type == "int4":
        packed = torch.empty((N, K // 8), dtype=torch.int32, device=tiled.device)
        scale = torch.empty((N, K // blocksize), dtype=torch.int32, device=tiled.device)
    else:
        packed = torch.empty((N * K // 2, 1), dtype=torch.uint8, device=tiled.device)
        scale = torch.empty((N * K // blocksize,), dtype=torch.float32, device=tiled.device)
    C.check(C.load().b200_untile_4bit(tiled.data_ptr(), N, K, QTYPES[quant_type], blocksize, packed.data_ptr(), scale.data_ptr(),
                                      C.stream_ptr(tiled.device)), "untile_4bit")
    return packed, scale


def qgemv_tiled(x: torch.Tensor, tiled: torch.Tensor, N: int, K: int, quant_type: str = "nf4", blocksize: int = 64, *,
                bias: Optional[torch.Tensor] = None, bias2: Optional[torch.Tensor] = None, delta: Optional[torch.Tensor] = None,
                x_out: Optional[torch.Tensor] = None, norm_weight: Optional[torch.Tensor] = None, norm_eps: float = 1e-5,
                residual: Optional[torch.Tensor] = None, tiled2: Optional[torch.Tensor] = None,
                prefetch: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None,
                tp: Optional["C.TpHandoff"] = None, fold: int = 1, kv_hint: Optional[tuple] = None,
                attn: Optional[dict] = None) -> torch.Tensor:
    """y = epilogue(prologue(x) @ dequant(W).T) over tiled matrices; see b200_qgemv_tiled in include/b200gpt.h.
    ``tp``: tensor-parallel hand-off over peer memory (lit_gpt_b200/p2p.py builds it).  ``fold``: the tiled buffer is
    the (N / fold, fold * K) view of the flat 4-bit storage (shards whose K is not a multiple of the block)."""
    C.require_cuda(x, tiled)
    x = _bf16c(x, "x")
    if x.shape[-1] != K:
        raise C.B200Error(f"qgemv_tiled: x has {x.shape[-1]} features, weight expects {K}")
    lead = x.shape[:-1]
    batch = x.numel() // K
    epilogue = C.EPI_SWIGLU if tiled2 is not None else (C.EPI_RESIDUAL if residual is not None else C.EPI_STORE)
    y = out if out is not None else torch.empty((*lead, N), dtype=torch.bfloat16, device=x.device)
    delta, residual, bias, bias2, norm_weight = (_bf16c(delta, "delta"), _bf16c(residual, "residual"), _bf16c(bias, "bias"),
                                                 _bf16c(bias2, "bias2"), _bf16c(norm_weight, "norm_weight"))
    ws = _ws("qgemv_tiled", x.device, C.load().b200_qgemv_tiled_workspace_bytes())
    hint = None
    if kv_hint is not None:  # (k_cache, v_cache, pos): L2 prefetch of the filled cache rows for the attention step that follows
        kc, vc, pos = kv_hint
        hint = C.KvHint(k=kc.data_ptr(), v=vc.data_ptr(), pos=pos.data_ptr(), pos_is_int64=int(pos.dtype == torch.int64),
                        chunks=kc.shape[0] * kc.shape[1], chunk_stride=kc.shape[2] * kc.shape[3] * kc.element_size(),
                        row_bytes=kc.shape[3] * kc.element_size(), max_rows=kc.shape[2])
    at = None
    if attn is not None:  # the decode-attention step fused into this (QKV) launch: y is its input, attn["out"] its output
        kc, vc, pos, cos, sin = attn["k_cache"], attn["v_cache"], attn["pos"], attn["cos"], attn["sin"]
        at = C.AttnDecodeArgs(
            qkv=y.data_ptr(), cos=cos.data_ptr(), sin=sin.data_ptr(), table_dtype=_table_code(cos), table_preselected=0,
            rope_n_elem=attn["rope_n_elem"], pos=pos.data_ptr(), pos_is_int64=int(pos.dtype == torch.int64), k_cache=kc.data_ptr(),
            v_cache=vc.data_ptr(), out=attn["out"].data_ptr(), batch=kc.shape[0], n_head=attn["n_head"], n_groups=attn["n_groups"],
            head_size=attn["head_size"], max_seq=kc.shape[2], num_splits=0, workspace=None, workspace_bytes=0)
    args = C.QGemvTiledArgs(
        tiled=tiled.data_ptr(), tiled2=C.ptr(tiled2), qtype=QTYPES[quant_type], blocksize=blocksize, N=N, K=K, x=x.data_ptr(),
        delta=C.ptr(delta), x_out=C.ptr(x_out), norm_weight=C.ptr(norm_weight), norm_eps=norm_eps, batch=batch, epilogue=epilogue,
        bias=C.ptr(bias), bias2=C.ptr(bias2), residual=C.ptr(residual), y=y.data_ptr(), prefetch=C.ptr(prefetch),
        prefetch_bytes=0 if prefetch is None else prefetch.numel() * prefetch.element_size(),
        workspace=ws.data_ptr(), workspace_bytes=ws.numel(), tp=None if tp is None else ctypes.pointer(tp), fold=fold,
        kv_hint=None if hint is None else ctypes.pointer(hint), attn=None if at is None else ctypes.addressof(at))
    C.check(C.load().b200_qgemv_tiled(ctypes.byref(args), C.stream_ptr(x.device)), "qgemv_tiled")
    return y


@functools.lru_cache(maxsize=None)
def qgemv_attn_fusable(N: int, K: int, quant_type: str, n_head: int, n_groups: int, head_size: int, rope_n_elem: int, max_seq: int) -> bool:
    """Whether b200_qgemv_tiled can run the decode attention of (n_head, n_groups, head_size) fused into the QKV GEMV (dry run)."""
    at = C.AttnDecodeArgs(qkv=256, cos=256, sin=256, table_dtype=0, table_preselected=0, rope_n_elem=rope_n_elem, pos=256,
                          pos_is_int64=1, k_cache=256, v_cache=256, out=256, batch=1, n_head=n_head, n_groups=n_groups,
                          head_size=head_size, max_seq=max_seq, num_splits=0, workspace=None, workspace_bytes=0)
    args = C.QGemvTiledArgs(qtype=QTYPES[quant_type], blocksize=128 if quant_type == "int4" else 64, N=N, K=K, batch=1,
                            epilogue=C.EPI_STORE, attn=ctypes.addressof(at))
    return bool(C.load().b200_qgemv_tiled_supported(ctypes.byref(args)))


@functools.lru_cache(maxsize=None)
def _qgemv_tiled_supported(N: int, K: int, quant_type: str, batch: int, swiglu: bool, fold: int) -> bool:
    args = C.QGemvTiledArgs(qtype=QTYPES[quant_type], blocksize=128 if quant_type == "int4" else 64, N=N, K=K, batch=batch,
                            epilogue=C.EPI_SWIGLU if swiglu else C.EPI_STORE, tiled2=256 if swiglu else None, fold=fold)  # dry run: pointers are not dereferenced
    return bool(C.load().b200_qgemv_tiled_supported(ctypes.byref(args)))


@torch._dynamo.assume_constant_result
def qgemv_tiled_supported(N: int, K: int, quant_type: str, batch: int, swiglu: bool = False, fold: int = 1) -> bool:
    """Whether b200_qgemv_tiled runs (batch, K) x (N, K) on a streaming kernel (its shared-memory plan covers the
    activation rows); asked up front so that the caller can pick the generic GEMV / the prefill GEMM instead.
    A pure function of the shape: a constant while tracing (no graph break)."""
    if not 1 <= batch <= 16:
        return False
    return _qgemv_tiled_supported(int(N), int(K), quant_type, int(batch), bool(swiglu), int(fold))


def tp_gather(h: "C.TpHandoff", batch: int, n_per_rank: int, device: torch.device) -> torch.Tensor:
    """All-gather of a row-sharded Linear's outputs over the peer-memory slots -> (batch, world * n_per_rank) bf16."""
    out = torch.empty((batch, h.world * n_per_rank), dtype=torch.bfloat16, device=device)
    ws = _ws("qgemv_tiled", device, C.load().b200_qgemv_tiled_workspace_bytes())
    C.check(C.load().b200_tp_gather(ctypes.byref(h), batch, n_per_rank, out.data_ptr(), ws.data_ptr(), C.stream_ptr(device)), "tp_gather")
    return out


def check_gemv_errors(device: torch.device) -> None:
    """Raise if a GEMV kernel on `device` gave up a bounded wait (a tensor-parallel peer that never delivered its
    packets, a stream-K partner CTA): synchronizes the current stream.  Called by generate() at its sync points."""
    key = ("qgemv_tiled", device.index if device.index is not None else torch.cuda.current_device())
    ws = _workspaces.get(key)
    if ws is None:
        return
    code = ctypes.c_uint(0)
    with torch.cuda.device(device):
        C.check(C.load().b200_qgemv_tiled_error(ws.data_ptr(), C.stream_ptr(device), ctypes.byref(code)), "qgemv_tiled_error")
    if code.value:
        what = {1: "a stream-K partner CTA", 2: "a tensor-parallel peer"}.get(code.value, f"code {code.value}")
        raise C.B200Error(f"dequant-GEMV: {what} did not deliver within the bounded wait; the output of this step is invalid")


def qgemv_tiled_tune(depth: int = 0, ctas_per_sm: int = 0, l2_prefetch: bool = True) -> None:
    C.load().b200_qgemv_tiled_tune(depth, ctas_per_sm, int(l2_prefetch))


# ---------------------------------------------------------------------------------------------------------
# norm / embedding / argmax
# ---------------------------------------------------------------------------------------------------------
def qgemm(x: torch.Tensor, packed: torch.Tensor, scale: torch.Tensor, N: int, K: int, quant_type: str = "nf4", blocksize: int = 64,
          bias: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Prefill GEMM y (M, N) = x (M, K) . dequant(W)^T on tcgen05 / TMEM (reference-layout packed weights + scales).
    Raises NotImplementedError for shapes the kernel does not cover (K not a multiple of the block, N % 8 != 0)."""
    C.require_cuda(x, packed, scale, bias)
    x = _bf16c(x, "x")
    bias = _bf16c(bias, "bias")
    M = x.numel() // K
    if x.numel() != M * K:
        raise C.B200Error("qgemm: x does not hold whole rows of length K")
    y = out if out is not None else torch.empty(tuple(x.shape[:-1]) + (N,), dtype=torch.bfloat16, device=x.device)
    need = C.load().b200_qgemm_workspace_bytes(M, N, K)   # fp32 partial sums of a one-tile prompt split over k ranges; 0 otherwise
    ws = _ws("qgemm", x.device, need) if need else None
    args = C.QGemmArgs(packed=packed.data_ptr(), scale=scale.data_ptr(), qtype=QTYPES[quant_type], blocksize=blocksize, M=M, N=N, K=K,
                       x=x.data_ptr(), bias=C.ptr(bias), y=y.data_ptr(), workspace=C.ptr(ws), workspace_bytes=0 if ws is None else ws.numel())
    C.check(C.load().b200_qgemm(ctypes.byref(args), C.stream_ptr(x.device)), "qgemm")
    return y


def qgemm_tune(mt3: bool = True) -> None:
    """development switch: 384-token tiles for long prompts (default) or the 256-token tiles of round 1"""
    C.check(C.load().b200_qgemm_tune(1 if mt3 else 0), "qgemm_tune")


def rmsnorm(x: torch.Tensor, weight: torch.Tensor, eps: float = 1e-5, delta: Optional[torch.Tensor] = None,
            sum_out: Optional[torch.Tensor] = None) -> torch.Tensor:
    C.require_cuda(x, weight)
    x, weight, delta = _bf16c(x, "x"), _bf16c(weight, "weight"), _bf16c(delta, "delta")
    dim = x.shape[-1]
    y = torch.empty_like(x)
    C.check(C.load().b200_rmsnorm(x.data_ptr(), C.ptr(delta), weight.data_ptr(), eps, y.data_ptr(), C.ptr(sum_out),
                                  x.numel() // dim, dim, C.stream_ptr(x.device)), "rmsnorm")
    return y


def swiglu(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """silu(a) * b with the reference's bf16 rounding points (model.py:715), one kernel."""
    C.require_cuda(a, b)
    a, b = _bf16c(a, "a"), _bf16c(b, "b")
    if a.shape != b.shape:
        raise C.B200Error("swiglu: shape mismatch")
    y = torch.empty_like(a)
    C.check(C.load().b200_swiglu(a.data_ptr(), b.data_ptr(), y.data_ptr(), a.numel(), C.stream_ptr(a.device)), "swiglu")
    return y


def embedding(idx: torch.Tensor, wte: torch.Tensor) -> torch.Tensor:
    C.require_cuda(idx, wte)
    wte = _bf16c(wte, "wte")
    if idx.dtype not in (torch.int64, torch.int32):
        raise C.B200Error(f"embedding: idx must be int32/int64 (got {idx.dtype})")
    idx = idx.contiguous()
    out = torch.empty((*idx.shape, wte.shape[1]), dtype=torch.bfloat16, device=wte.device)
    C.check(C.load().b200_embedding(idx.data_ptr(), int(idx.dtype == torch.int64), wte.data_ptr(), out.data_ptr(), idx.numel(),
                                    wte.shape[1], wte.shape[0], C.stream_ptr(wte.device)), "embedding")
    return out


def argmax_(logits: torch.Tensor, token_out: torch.Tensor, pos_inout: Optional[torch.Tensor] = None, vocab: Optional[int] = None,
            history: Optional[torch.Tensor] = None) -> None:
    """token_out[b] = argmax(logits[b, :vocab]); optionally pos_inout[0] += 1 and history[pos] = token (on device)."""
    if history is not None and history.dtype != token_out.dtype:
        raise C.B200Error("argmax_: history must have the dtype of token_out")
    C.require_cuda(logits, token_out)
    logits = _bf16c(logits, "logits")
    V = logits.shape[-1]
    batch = logits.numel() // V
    C.check(C.load().b200_argmax(logits.data_ptr(), batch, vocab or V, V, token_out.data_ptr(), int(token_out.dtype == torch.int64),
                                 C.ptr(pos_inout), int(pos_inout is not None and pos_inout.dtype == torch.int64),
                                 C.ptr(history), 0 if history is None else history.numel(),
                                 C.stream_ptr(logits.device)), "argmax")


def new_rng_state(device: torch.device, seed: Optional[int] = None) -> torch.Tensor:
    """Device-side Philox state for ``sample_``: int64[2] = (seed, offset).  The seed defaults to torch's current seed, so
    ``seed_everything`` (generate/base.py:140) makes the draws reproducible."""
    seed = (torch.initial_seed() if seed is None else int(seed)) & 0x7FFFFFFFFFFFFFFF  # torch's default seed is a full uint64
    return torch.tensor([seed, 0], dtype=torch.int64, device=device)


def sample_(logits: torch.Tensor, token_out: torch.Tensor, temperature: float = 1.0, top_k: Optional[int] = None,
            rng_state: Optional[torch.Tensor] = None, pos_inout: Optional[torch.Tensor] = None, history: Optional[torch.Tensor] = None,
            noise: Optional[torch.Tensor] = None) -> None:
    """token_out[0] = one draw from softmax(top_k(logits[0, -1]) / temperature) (generate/base.py:22-41), on the device;
    optionally pos_inout[0] += 1 and history[pos] = token like ``argmax_``.  ``noise``: Exp(1) values to use instead of Philox."""
    C.require_cuda(logits, token_out)
    if logits.dtype not in (torch.bfloat16, torch.float32):
        raise C.B200Error(f"sample_: logits must be bf16 or fp32 (got {logits.dtype})")
    V = logits.shape[-1]
    row = (logits[0, -1] if logits.dim() == 3 else logits[0] if logits.dim() == 2 else logits).contiguous()  # base.py:31
    if history is not None and history.dtype != token_out.dtype:
        raise C.B200Error("sample_: history must have the dtype of token_out")
    if temperature > 0.0 and rng_state is None and noise is None:
        raise C.B200Error("sample_: temperature > 0 needs rng_state (K.new_rng_state) or noise")
    if rng_state is not None and (rng_state.dtype != torch.int64 or rng_state.numel() != 2 or not rng_state.is_cuda):
        raise C.B200Error("sample_: rng_state must be a CUDA int64[2] tensor")
    if noise is not None:
        noise = noise.to(torch.float32).contiguous()
        if noise.numel() != V:
            raise C.B200Error("sample_: noise must hold one value per vocabulary entry")
    C.check(C.load().b200_sample(row.data_ptr(), C.DTYPE_CODE[row.dtype], V, float(temperature), int(top_k or 0), C.ptr(rng_state), C.ptr(noise),
                                 token_out.data_ptr(), int(token_out.dtype == torch.int64), C.ptr(pos_inout),
                                 int(pos_inout is not None and pos_inout.dtype == torch.int64), C.ptr(history),
                                 0 if history is None else history.numel(), C.stream_ptr(logits.device)), "sample")


# ---------------------------------------------------------------------------------------------------------
# attention
# ---------------------------------------------------------------------------------------------------------
def _table_code(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return C.F32
    if t.dtype == torch.bfloat16:
        return C.BF16
    raise C.B200Error(f"cos/sin tables must be fp32 or bf16 (got {t.dtype})")


def attn_decode(qkv: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, pos: torch.Tensor, k_cache: torch.Tensor,
                v_cache: torch.Tensor, n_head: int, n_groups: int, head_size: int, rope_n_elem: int,
                num_splits: int = 0, table_preselected: bool = False) -> torch.Tensor:
    """Fused RoPE + cache append + split-K decode attention for one token per sequence.

    qkv (B, 1, (H+2G)*hs) or (B, (H+2G)*hs); cos/sin: the model's FULL tables (max_seq, rope_n_elem); pos: device int
    scalar tensor (input_pos); caches (B, G, max_seq, hs) bf16, updated in place.  Returns (B, H*hs)."""
    C.require_cuda(qkv, cos, sin, pos, k_cache, v_cache)
    qkv = _bf16c(qkv, "qkv")
    B = k_cache.shape[0]
    max_seq = k_cache.shape[2]
    if k_cache.dtype != torch.bfloat16 or not k_cache.is_contiguous() or not v_cache.is_contiguous():
        raise C.B200Error("kv cache must be contiguous bfloat16 (B, n_query_groups, max_seq, head_size)")
    if qkv.numel() != B * (n_head + 2 * n_groups) * head_size:
        raise C.B200Error("attn_decode: qkv does not hold exactly one token per cached sequence")
    if pos.dtype not in (torch.int64, torch.int32) or pos.numel() != 1:
        raise C.B200Error("attn_decode: pos must be a single int32/int64 device scalar")
    cos, sin = cos.contiguous(), sin.contiguous()
    out = torch.empty((B, n_head * head_size), dtype=torch.bfloat16, device=qkv.device)
    lib = C.load()
    max_splits = 64
    ws = _ws("attn", qkv.device, lib.b200_attn_decode_workspace_bytes(B, n_head, head_size, max_splits))
    args = C.AttnDecodeArgs(
        qkv=qkv.data_ptr(), cos=cos.data_ptr(), sin=sin.data_ptr(), table_dtype=_table_code(cos), table_preselected=int(table_preselected), rope_n_elem=rope_n_elem,
        pos=pos.data_ptr(), pos_is_int64=int(pos.dtype == torch.int64), k_cache=k_cache.data_ptr(), v_cache=v_cache.data_ptr(),
        out=out.data_ptr(), batch=B, n_head=n_head, n_groups=n_groups, head_size=head_size, max_seq=max_seq,
        num_splits=num_splits, workspace=ws.data_ptr(), workspace_bytes=ws.numel())
    C.check(lib.b200_attn_decode(ctypes.byref(args), C.stream_ptr(qkv.device)), "attn_decode")
    return out


def rope_kv_append(qkv: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, pos: torch.Tensor, k_cache: torch.Tensor,
                   v_cache: torch.Tensor, n_head: int, n_groups: int, head_size: int, rope_n_elem: int,
                   table_preselected: bool = False) -> torch.Tensor:
    """Prefill: RoPE q/k for T tokens at positions pos[0..T), write k/v into the GQA-native caches, return roped q
    as (B, H, T, hs)."""
    C.require_cuda(qkv, cos, sin, pos, k_cache, v_cache)
    qkv = _bf16c(qkv, "qkv")
    B, T = qkv.shape[0], qkv.shape[1]
    if pos.numel() != T:
        raise C.B200Error("rope_kv_append: pos must have one entry per token")
    cos, sin, pos = cos.contiguous(), sin.contiguous(), pos.contiguous()
    q = torch.empty((B, n_head, T, head_size), dtype=torch.bfloat16, device=qkv.device)
    C.check(C.load().b200_rope_kv_append(qkv.data_ptr(), cos.data_ptr(), sin.data_ptr(), _table_code(cos), int(table_preselected), rope_n_elem,
                                         pos.data_ptr(), int(pos.dtype == torch.int64), k_cache.data_ptr(), v_cache.data_ptr(),
                                         q.data_ptr(), B, T, n_head, n_groups, head_size, k_cache.shape[2],
                                         C.stream_ptr(qkv.device)), "rope_kv_append")
    return q


def attn_prefill(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, pos: torch.Tensor, scale: Optional[float] = None) -> torch.Tensor:
    """Prefill attention (reference ``scaled_dot_product_attention``, lit_gpt/model.py:658-665, cached call): q (B, H, T, hs)
    roped, caches (B, G, Tmax, hs) already extended; row t sees cache positions <= pos[t].  Returns (B, T, H * hs)."""
    C.require_cuda(q, k_cache, v_cache, pos)
    q = _bf16c(q, "q")
    if k_cache.dtype != torch.bfloat16 or v_cache.dtype != torch.bfloat16 or not (k_cache.is_contiguous() and v_cache.is_contiguous()):
        raise C.B200Error("attn_prefill: caches must be contiguous bf16")
    B, H, T, hs = q.shape
    G, Tmax = k_cache.shape[1], k_cache.shape[2]
    if pos.numel() != T:
        raise C.B200Error("attn_prefill: pos must have one entry per token")
    pos = pos.contiguous()
    y = torch.empty((B, T, H * hs), dtype=torch.bfloat16, device=q.device)
    C.check(C.load().b200_attn_prefill(q.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(), pos.data_ptr(), int(pos.dtype == torch.int64),
                                       y.data_ptr(), B, T, H, G, hs, Tmax, float(scale if scale is not None else hs ** -0.5),
                                       C.stream_ptr(q.device)), "attn_prefill")
    return y


def attn_decode_tune(force_scalar: bool = False) -> bool:
    """Test knob: route grouped-query shapes through the scalar decode-attention kernel; returns the previous setting."""
    return bool(C.load().b200_attn_decode_tune(int(force_scalar)))


def set_pdl(enabled: bool) -> bool:
    return bool(C.load().b200_set_pdl(int(enabled)))


def launch_count() -> int:
    return int(C.load().b200_launch_count())


def qgemv_tune(ksteps: int = 0, ctas_per_sm: int = 0, force_generic: bool = False) -> None:
    C.load().b200_qgemv_tune(ksteps, ctas_per_sm, int(force_generic))

