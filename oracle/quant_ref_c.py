"""ctypes binding of oracle/quant_ref.c (TEST INFRASTRUCTURE; see oracle/__init__.py)."""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libquant_ref.so")
_lib = None
_QT = {"nf4": 0, "fp4": 1}


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "quant_ref.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        # -march=native objects must not travel between hosts with different CPUs: rebuild when possible
        subprocess.check_call(["make", "-C", _HERE, "-B", "-s"])
    return _SO


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        try:
            build()
        except Exception:
            if not os.path.exists(_SO):
                raise
        _lib = ctypes.CDLL(_SO)
        _lib.ref_num_threads.restype = ctypes.c_int
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def num_threads() -> int:
    return int(lib().ref_num_threads())


def quantize_4bit(w: np.ndarray, quant_type="nf4", blocksize=64):
    flat = np.ascontiguousarray(w, np.float32).reshape(-1)
    n = flat.size
    packed = np.zeros((n + 1) // 2, np.uint8)
    absmax = np.zeros((n + blocksize - 1) // blocksize, np.float32)
    lib().ref_quantize_4bit(_p(flat), ctypes.c_int64(n), blocksize, _QT[quant_type], _p(packed), _p(absmax))
    return packed.reshape(-1, 1), absmax


def dequantize_4bit(packed, absmax, shape, quant_type="nf4", blocksize=64):
    n = int(np.prod(shape))
    out = np.empty(n, np.float32)
    packed = np.ascontiguousarray(packed).reshape(-1)
    absmax = np.ascontiguousarray(absmax, np.float32)
    lib().ref_dequantize_4bit(_p(packed), _p(absmax), ctypes.c_int64(n), blocksize, _QT[quant_type], _p(out))
    return out.reshape(shape)


def gemv_4bit(packed, absmax, x, N, K, quant_type="nf4", blocksize=64, round_w_bf16=True):
    x = np.ascontiguousarray(x, np.float32).reshape(-1, K)
    B = x.shape[0]
    y = np.empty((B, N), np.float32)
    packed = np.ascontiguousarray(packed).reshape(-1)
    absmax = np.ascontiguousarray(absmax, np.float32)
    lib().ref_gemv_4bit(_p(packed), _p(absmax), _p(x), _p(y), ctypes.c_int64(N), ctypes.c_int64(K), B, blocksize,
                        _QT[quant_type], int(round_w_bf16))
    return y


def dequantize_int4(qweight, scales, qzeros, K, group=128):
    N = qweight.shape[0]
    out = np.empty((N, K), np.float32)
    qweight = np.ascontiguousarray(qweight, np.uint32)
    scales = np.ascontiguousarray(scales, np.float16)
    qzeros = np.ascontiguousarray(qzeros, np.uint32)
    lib().ref_dequantize_int4(_p(qweight), _p(scales.view(np.uint16)), _p(qzeros), ctypes.c_int64(N),
                              ctypes.c_int64(K), group, _p(out))
    return out


def gemv_int4(qweight, scales, qzeros, x, K, group=128, round_w_bf16=True):
    N = qweight.shape[0]
    x = np.ascontiguousarray(x, np.float32).reshape(-1, K)
    B = x.shape[0]
    y = np.empty((B, N), np.float32)
    qweight = np.ascontiguousarray(qweight, np.uint32)
    scales = np.ascontiguousarray(scales, np.float16)
    qzeros = np.ascontiguousarray(qzeros, np.uint32)
    lib().ref_gemv_int4(_p(qweight), _p(scales.view(np.uint16)), _p(qzeros), _p(x), _p(y), ctypes.c_int64(N),
                        ctypes.c_int64(K), B, group, int(round_w_bf16))
    return y
