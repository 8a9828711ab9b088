"""ctypes binding of ``libb200gpt.so`` — the ONLY way the Python package reaches the kernels.

There is no fallback: if the library is missing and cannot be built, importing any kernel raises.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_int64, c_size_t, c_void_p

import torch

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "lib", "libb200gpt.so")

BF16, F16, F32 = 0, 1, 2
Q_NF4, Q_FP4, Q_INT4 = 0, 1, 2
EPI_STORE, EPI_RESIDUAL, EPI_SWIGLU = 0, 1, 2
DTYPE_CODE = {torch.bfloat16: BF16, torch.float16: F16, torch.float32: F32}


class LinearInt8Args(ctypes.Structure):  # mirrors b200_linear_int8_args
    _fields_ = [("x", c_void_p), ("cb", c_void_p), ("scb", c_void_p), ("bias", c_void_p), ("y", c_void_p),
                ("M", c_int64), ("N", c_int64), ("K", c_int64), ("threshold", c_float), ("workspace", c_void_p),
                ("workspace_bytes", c_size_t)]


class QGemvArgs(ctypes.Structure):
    _fields_ = [
        ("packed", c_void_p), ("scale", c_void_p), ("packed2", c_void_p), ("scale2", c_void_p),
        ("qtype", c_int), ("blocksize", c_int), ("N", c_int64), ("K", c_int64),
        ("x", c_void_p), ("delta", c_void_p), ("x_out", c_void_p), ("norm_weight", c_void_p),
        ("norm_eps", c_float), ("batch", c_int), ("epilogue", c_int),
        ("bias", c_void_p), ("residual", c_void_p), ("y", c_void_p),
        ("workspace", c_void_p), ("workspace_bytes", c_size_t),
    ]


class TpHandoff(ctypes.Structure):  # mirrors b200_tp_handoff
    _fields_ = [("world", c_int), ("rank", c_int), ("peer_y", c_void_p * 8), ("reduced", c_void_p), ("reserved", c_void_p * 7), ("send_seq", c_void_p),
                ("slots", c_void_p), ("slot_stride", c_int64), ("recv_seq", c_void_p), ("recv_done", c_void_p)]


class KvHint(ctypes.Structure):
    _fields_ = [("k", c_void_p), ("v", c_void_p), ("pos", c_void_p), ("pos_is_int64", c_int), ("chunks", c_int),
                ("chunk_stride", c_int64), ("row_bytes", c_int), ("max_rows", c_int)]


class QGemvTiledArgs(ctypes.Structure):
    _fields_ = [
        ("tiled", c_void_p), ("tiled2", c_void_p), ("qtype", c_int), ("blocksize", c_int), ("N", c_int64), ("K", c_int64),
        ("x", c_void_p), ("delta", c_void_p), ("x_out", c_void_p), ("norm_weight", c_void_p),
        ("norm_eps", c_float), ("batch", c_int), ("epilogue", c_int),
        ("bias", c_void_p), ("bias2", c_void_p), ("residual", c_void_p), ("y", c_void_p),
        ("prefetch", c_void_p), ("prefetch_bytes", c_int64), ("workspace", c_void_p), ("workspace_bytes", c_size_t),
        ("tp", ctypes.POINTER(TpHandoff)), ("kv_hint", ctypes.POINTER(KvHint)), ("attn", c_void_p), ("fold", c_int),
    ]


class AttnDecodeArgs(ctypes.Structure):
    _fields_ = [
        ("qkv", c_void_p), ("cos", c_void_p), ("sin", c_void_p), ("table_dtype", c_int), ("table_preselected", c_int), ("rope_n_elem", c_int),
        ("pos", c_void_p), ("pos_is_int64", c_int), ("k_cache", c_void_p), ("v_cache", c_void_p), ("out", c_void_p),
        ("batch", c_int), ("n_head", c_int), ("n_groups", c_int), ("head_size", c_int), ("max_seq", c_int),
        ("num_splits", c_int), ("workspace", c_void_p), ("workspace_bytes", c_size_t),
    ]


class QGemmArgs(ctypes.Structure):
    _fields_ = [
        ("packed", c_void_p), ("scale", c_void_p), ("qtype", c_int), ("blocksize", c_int),
        ("M", c_int64), ("N", c_int64), ("K", c_int64), ("x", c_void_p), ("bias", c_void_p), ("y", c_void_p),
        ("workspace", c_void_p), ("workspace_bytes", c_size_t),
    ]


_SIGNATURES = {
    "b200_abi_version": (c_int, []),
    "b200_last_error": (c_char_p, []),
    "b200_device_info": (c_int, [ctypes.POINTER(c_int)] * 3),
    "b200_set_pdl": (c_int, [c_int]),
    "b200_launch_count": (c_int64, []),
    "b200_qgemv_tune": (c_int, [c_int, c_int, c_int]),
    "b200_quantize_4bit": (c_int, [c_void_p, c_int, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "b200_dequantize_4bit": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p, c_int, c_void_p]),
    "b200_double_quant_absmax": (c_int, [c_void_p, c_int64, c_void_p, c_float, c_void_p, c_void_p, c_void_p, c_void_p]),
    "b200_p2p_alloc": (c_int, [c_size_t, ctypes.POINTER(c_void_p), c_void_p]),
    "b200_p2p_open": (c_int, [c_void_p, ctypes.POINTER(c_void_p)]),
    "b200_p2p_close": (c_int, [c_void_p]),
    "b200_p2p_free": (c_int, [c_void_p]),
    "b200_quantize_int8": (c_int, [c_void_p, c_int, c_int64, c_int64, c_void_p, c_void_p, c_void_p]),
    "b200_dequantize_int8": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_int, c_void_p]),
    "b200_linear_int8_workspace_bytes": (c_size_t, [c_int64, c_int64]),
    "b200_linear_int8": (c_int, [ctypes.POINTER(LinearInt8Args), c_void_p]),
    "b200_quantize_int4": (c_int, [c_void_p, c_int, c_int64, c_int64, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "b200_dequantize_int4": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int, c_void_p, c_int, c_void_p]),
    "b200_qgemv_workspace_bytes": (c_size_t, []),
    "b200_qgemv": (c_int, [ctypes.POINTER(QGemvArgs), c_void_p]),
    "b200_tiled_bytes": (c_size_t, [c_int64, c_int64, c_int]),
    "b200_tile_4bit": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int, c_int, c_void_p, c_void_p]),
    "b200_untile_4bit": (c_int, [c_void_p, c_int64, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "b200_qgemv_tiled": (c_int, [ctypes.POINTER(QGemvTiledArgs), c_void_p]),
    "b200_qgemv_tiled_workspace_bytes": (c_size_t, []),
    "b200_qgemv_tiled_tune": (c_int, [c_int, c_int, c_int]),
    "b200_qgemv_tiled_trace": (c_int, [c_void_p, c_int]),
    "b200_qgemv_tiled_supported": (c_int, [ctypes.POINTER(QGemvTiledArgs)]),
    "b200_qgemv_tiled_error": (c_int, [c_void_p, c_void_p, ctypes.POINTER(ctypes.c_uint)]),
    "b200_rmsnorm": (c_int, [c_void_p, c_void_p, c_void_p, c_float, c_void_p, c_void_p, c_int64, c_int64, c_void_p]),
    "b200_attn_decode_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int]),
    "b200_attn_decode": (c_int, [ctypes.POINTER(AttnDecodeArgs), c_void_p]),
    "b200_attn_decode_tune": (c_int, [c_int]),
    "b200_attn_decode_noalloc": (c_int, [ctypes.POINTER(AttnDecodeArgs), c_void_p]),
    "b200_attn_decode_noalloc_max_bytes": (c_size_t, []),
    "b200_rope_kv_append": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p,
                                    c_int, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "b200_attn_prefill": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int,
                                  c_float, c_void_p]),
    "b200_tp_gather": (c_int, [ctypes.POINTER(TpHandoff), c_int, c_int64, c_void_p, c_void_p, c_void_p]),
    "b200_swiglu": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    "b200_embedding": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_void_p]),
    "b200_argmax": (c_int, [c_void_p, c_int64, c_int64, c_int64, c_void_p, c_int, c_void_p, c_int, c_void_p, c_int64, c_void_p]),
    "b200_sample": (c_int, [c_void_p, c_int, c_int64, c_float, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p, c_int64,
                            c_void_p]),
    "b200_qgemm_workspace_bytes": (c_size_t, [c_int64, c_int64, c_int64]),
    "b200_qgemm": (c_int, [ctypes.POINTER(QGemmArgs), c_void_p]),
    "b200_qgemm_tune": (c_int, [c_int]),
}

_lib = None


class B200Error(RuntimeError):
    pass


def load(build_if_missing: bool = True) -> ctypes.CDLL:
    """Load libb200gpt.so (building it in-tree with nvcc when absent).  Raises if neither is possible."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        if not build_if_missing:
            raise B200Error(f"{LIB_PATH} is missing: run `python -m lit_gpt_b200.build`")
        from .build import build_library

        build_library()
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here = header/library mismatch, which must be loud
        fn.restype = res
        fn.argtypes = args
    if lib.b200_abi_version() != 1:
        raise B200Error(f"libb200gpt ABI version {lib.b200_abi_version()} != 1")
    _lib = lib
    return lib


def exported_symbols() -> list[str]:
    return sorted(_SIGNATURES)


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().b200_last_error().decode(errors="replace")
        if rc == 3:
            raise NotImplementedError(f"{what}: {msg}")
        raise B200Error(f"{what} failed (code {rc}): {msg}")


def ptr(t: torch.Tensor | None) -> int | None:
    return None if t is None else t.data_ptr()


def stream_ptr(device: torch.device | None = None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def require_cuda(*tensors: torch.Tensor | None) -> None:
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise B200Error("b200gpt kernels run on CUDA tensors only; there is no CPU fallback")
