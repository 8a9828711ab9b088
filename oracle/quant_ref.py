"""CPU restatement (numpy) of the weight-only quantisation formats on the hot path.

TEST INFRASTRUCTURE (see oracle/__init__.py) — never imported by the product.

PARITY UNPINNED for every function in this file: the arithmetic of
``--quantize bnb.nf4 / bnb.nf4-dq / bnb.fp4 / bnb.int8`` (reference call sites
generate/base.py:105,129-134; generate/tp.py:104,127-134; lit_gpt/lora.py:156-160;
lit_gpt/utils.py:85-87) lives in ``bitsandbytes==0.41.0``
(requirements-all.txt:3), reached through ``lightning@ed367ca``
(requirements.txt:2).  Neither is vendored under /root/reference nor installable
here, and the reference's own tests hold no golden vectors for it.
What follows restates the published algorithm of bitsandbytes 0.41.x
(csrc/kernels.cu: kQuantizeBlockwise / kDequantizeBlockwise / dQuantizeNF4 /
dQuantizeFP4 / dDequantizeFP4Tree / dQuantize<0>; bitsandbytes/functional.py:
quantize_4bit / dequantize_4bit / create_dynamic_map / quantize_blockwise).
``int4 group-128`` is this project's own format (the reference has none).
"""
from __future__ import annotations

import numpy as np

F32 = np.float32

# ---------------------------------------------------------------------------------------------
# 4-bit codebooks
# ---------------------------------------------------------------------------------------------
# bitsandbytes functional.py get_4bit_type("nf4") / kernels.cu dDequantizeNF4
NF4_CODE = np.array(
    [
        -1.0,
        -0.6961928009986877,
        -0.5250730514526367,
        -0.39491748809814453,
        -0.28444138169288635,
        -0.18477343022823334,
        -0.09105003625154495,
        0.0,
        0.07958029955625534,
        0.16093020141124725,
        0.24611230194568634,
        0.33791524171829224,
        0.44070982933044434,
        0.5626170039176941,
        0.7229568362236023,
        1.0,
    ],
    dtype=F32,
)
# kernels.cu dQuantizeNF4: decision tree of `x > t` tests; equals "count thresholds below x".
NF4_THRESHOLDS = np.array(
    [
        -0.8480964004993439,
        -0.6106329262256622,
        -0.4599952697753906,
        -0.33967943489551544,
        -0.23460740596055984,
        -0.13791173323988914,
        -0.045525018125772476,
        0.03979014977812767,
        0.1202552504837513,
        0.2035212516784668,
        0.2920137718319893,
        0.3893125355243683,
        0.5016634166240692,
        0.6427869200706482,
        0.8614784181118011,
    ],
    dtype=F32,
)

# kernels.cu dDequantizeFP4Tree: sign bit 3, magnitude by 3-bit code
_FP4_MAG = np.array(
    [0.0, 5.208333333e-03, 0.66666667, 1.0, 0.33333333, 0.5, 0.16666667, 0.25], dtype=F32
)
FP4_CODE = np.concatenate([_FP4_MAG, -_FP4_MAG]).astype(F32)
# kernels.cu dQuantizeFP4: thresholds on |x| (ascending) and the code each magnitude rank maps to
FP4_THRESHOLDS = np.array(
    [0.00260417, 0.0859375, 0.20833333, 0.29166667, 0.4166667, 0.583333, 0.8333333], dtype=F32
)
FP4_RANK_TO_CODE = np.array([0b000, 0b001, 0b110, 0b111, 0b100, 0b101, 0b010, 0b011], dtype=np.uint8)

CODEBOOKS = {"nf4": NF4_CODE, "fp4": FP4_CODE}


def _encode_nf4(xn: np.ndarray) -> np.ndarray:
    # NaN compares false everywhere -> code 0, as in the tree
    with np.errstate(invalid="ignore"):
        return (xn[..., None] > NF4_THRESHOLDS).sum(-1).astype(np.uint8)


def _encode_fp4(xn: np.ndarray) -> np.ndarray:
    with np.errstate(invalid="ignore"):
        sign = np.where(xn < 0, 8, 0).astype(np.uint8)
        rank = (np.abs(xn)[..., None] > FP4_THRESHOLDS).sum(-1)
    return (FP4_RANK_TO_CODE[rank] + sign).astype(np.uint8)


def quantize_4bit(w: np.ndarray, quant_type: str = "nf4", blocksize: int = 64):
    """functional.quantize_4bit + kQuantizeBlockwise<.., NF4|FP4>.

    w: any shape, float32 values (a bf16/fp16 weight is passed as its exact float32 image).
    Flattened row-major; per `blocksize` run: absmax (fp32); x * (1.0f/absmax) in fp32; tree;
    two codes per byte, EVEN element in the HIGH nibble.  Returns (packed uint8 [(n+1)//2, 1], absmax fp32 [nblocks]).
    """
    flat = np.ascontiguousarray(w, dtype=F32).reshape(-1)
    n = flat.size
    nblocks = (n + blocksize - 1) // blocksize
    pad = nblocks * blocksize - n
    padded = np.concatenate([flat, np.zeros(pad, F32)]) if pad else flat
    blocks = padded.reshape(nblocks, blocksize)
    absmax = np.abs(blocks).max(axis=1).astype(F32)
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = (F32(1.0) / absmax).astype(F32)
        xn = (blocks * inv[:, None]).astype(F32)
    codes = (_encode_nf4(xn) if quant_type == "nf4" else _encode_fp4(xn)).reshape(-1)[:n]
    if n % 2:
        codes = np.concatenate([codes, np.zeros(1, np.uint8)])
    packed = ((codes[0::2] << 4) | codes[1::2]).astype(np.uint8).reshape(-1, 1)
    return packed, absmax


def unpack_codes(packed: np.ndarray, n: int) -> np.ndarray:
    p = packed.reshape(-1)
    codes = np.empty(p.size * 2, np.uint8)
    codes[0::2] = p >> 4
    codes[1::2] = p & 0x0F
    return codes[:n]


def dequantize_4bit(packed, absmax, shape, quant_type: str = "nf4", blocksize: int = 64) -> np.ndarray:
    """kDequantizeBlockwise<.., NF4|FP4>: code[c] * absmax[block] in fp32 (caller casts to the compute dtype)."""
    n = int(np.prod(shape))
    codes = unpack_codes(packed, n)
    vals = CODEBOOKS[quant_type][codes]
    scale = np.repeat(np.asarray(absmax, F32), blocksize)[:n]
    return (vals * scale).astype(F32).reshape(shape)


# ---------------------------------------------------------------------------------------------
# double quantisation of absmax ("nf4-dq" / "fp4-dq"): functional.quantize_4bit(compress_statistics=True)
# ---------------------------------------------------------------------------------------------
def create_dynamic_map(signed: bool = True, max_exponent_bits: int = 7, total_bits: int = 8) -> np.ndarray:
    """functional.create_dynamic_map — the 256-entry 8-bit "dynamic" code, sorted ascending."""
    data = []
    non_sign_bits = total_bits - (1 if signed else 0)
    additional_items = 2 ** (non_sign_bits - max_exponent_bits) - 1
    if not signed:
        additional_items = 2 * additional_items
    i = 0
    for i in range(max_exponent_bits):
        fraction_items = int(
            2 ** (i + non_sign_bits - max_exponent_bits) + 1
            if signed
            else 2 ** (i + non_sign_bits - max_exponent_bits + 1) + 1
        )
        boundaries = np.linspace(0.1, 1, fraction_items, dtype=F32)
        means = ((boundaries[:-1] + boundaries[1:]) / F32(2.0)).astype(F32)
        scale = 10 ** (-(max_exponent_bits - 1) + i)
        data += (scale * means.astype(np.float64)).tolist()
        if signed:
            data += (-(scale) * means.astype(np.float64)).tolist()
    if additional_items > 0:
        boundaries = np.linspace(0.1, 1, additional_items + 1, dtype=F32)
        means = ((boundaries[:-1] + boundaries[1:]) / F32(2.0)).astype(F32)
        scale = 10 ** (-(max_exponent_bits - 1) + i)
        data += (scale * means.astype(np.float64)).tolist()
        if signed:
            data += (-(scale) * means.astype(np.float64)).tolist()
    data.append(0)
    data.append(1.0)
    data += [0] * (256 - len(data))
    data.sort()
    return np.array(data, dtype=F32)


DYNAMIC_MAP = create_dynamic_map()


def _dquantize8(code: np.ndarray, x: np.ndarray) -> np.ndarray:
    """kernels.cu dQuantize<0>(smem_code, 0.0f, x): pivot search then nearest of {lower, pivot, upper}."""
    out = np.empty(x.shape, np.uint8)
    flat_x = x.reshape(-1)
    flat_o = out.reshape(-1)
    for n, xv in enumerate(flat_x):
        pivot, upper_pivot, lower_pivot = 127, 255, 0
        lower, upper = F32(-1.0), F32(1.0)
        val = code[pivot]
        i = 64
        while i > 0:
            if xv > val:
                lower_pivot, lower = pivot, val
                pivot += i
            else:
                upper_pivot, upper = pivot, val
                pivot -= i
            val = code[pivot]
            i >>= 1
        if upper_pivot == 255:
            upper = code[upper_pivot]
        if lower_pivot == 0:
            lower = code[lower_pivot]
        if xv > val:
            midpoint = F32((upper + val) * F32(0.5))
            flat_o[n] = upper_pivot if xv > midpoint else pivot
        else:
            midpoint = F32((lower + val) * F32(0.5))
            flat_o[n] = lower_pivot if xv < midpoint else pivot
    return out


def quantize_blockwise8(a: np.ndarray, blocksize: int = 256, code: np.ndarray = DYNAMIC_MAP):
    """functional.quantize_blockwise + kQuantizeBlockwise<.., General8bit>."""
    flat = np.asarray(a, F32).reshape(-1)
    n = flat.size
    nblocks = (n + blocksize - 1) // blocksize
    q = np.zeros(n, np.uint8)
    absmax2 = np.zeros(nblocks, F32)
    for b in range(nblocks):
        seg = flat[b * blocksize : (b + 1) * blocksize]
        am = F32(np.abs(seg).max())
        absmax2[b] = am
        with np.errstate(divide="ignore", invalid="ignore"):
            xn = (seg * (F32(1.0) / am)).astype(F32)
        q[b * blocksize : b * blocksize + seg.size] = _dquantize8(code, xn)
    return q, absmax2


def dequantize_blockwise8(q, absmax2, blocksize: int = 256, code: np.ndarray = DYNAMIC_MAP) -> np.ndarray:
    scale = np.repeat(np.asarray(absmax2, F32), blocksize)[: q.size]
    return (code[q] * scale).astype(F32)


def double_quantize_absmax(absmax: np.ndarray, offset=None):
    """quantize_4bit(compress_statistics=True): offset = mean; absmax -= offset; 8-bit blockwise (256).
    `offset` may be given (a GPU mean reduces in a different order than numpy's; the rest is order-independent)."""
    absmax = np.asarray(absmax, F32)
    offset = F32(absmax.mean(dtype=F32)) if offset is None else F32(offset)
    q, absmax2 = quantize_blockwise8((absmax - offset).astype(F32), 256)
    return q, absmax2, offset


def dequantize_absmax(q, absmax2, offset) -> np.ndarray:
    """dequantize_4bit with nested state: dequantize_blockwise(qabsmax) + offset."""
    return (dequantize_blockwise8(q, absmax2, 256) + F32(offset)).astype(F32)


# ---------------------------------------------------------------------------------------------
# int4 group-128 (this project's format; SURVEY F2: the reference has no gptq/int4 code)
# ---------------------------------------------------------------------------------------------
def _f16(x):
    return np.asarray(x, np.float16)


def quantize_int4_g128(w: np.ndarray, group: int = 128):
    """Asymmetric round-to-nearest int4, groups of `group` along in_features, GPTQ storage convention.

    w: (N, K) float32.  Per (row, group): lo=min(w,0), hi=max(w,0); scale=fp16((hi-lo)/15) (1.0 if the
    group is all zero); zero=clamp(rint(-lo/scale),0,15); q=clamp(rint(w/scale)+zero,0,15), all in fp32
    with the fp16-rounded scale.  Returns qweight uint32 (N, ceil(K/8)) with k=8i+j in bits [4j,4j+4),
    scales fp16 (N, G), qzeros uint32 (N, ceil(G/8)) packed the same way, G=ceil(K/group).
    """
    w = np.asarray(w, F32)
    N, K = w.shape
    G = (K + group - 1) // group
    Kp = G * group
    wp = np.zeros((N, Kp), F32)
    wp[:, :K] = w
    g = wp.reshape(N, G, group)
    lo = np.minimum(g.min(-1), 0).astype(F32)
    hi = np.maximum(g.max(-1), 0).astype(F32)
    scale = ((hi - lo) / F32(15.0)).astype(F32)
    scale = np.where(scale == 0, F32(1.0), scale)
    scale16 = _f16(scale)
    scale16 = np.where(scale16 == 0, np.float16(6.1035e-05), scale16)  # fp16 underflow guard
    s = scale16.astype(F32)
    zero = np.clip(np.rint((-lo) / s), 0, 15).astype(F32)
    q = np.clip(np.rint(g / s[..., None]) + zero[..., None], 0, 15).astype(np.uint32).reshape(N, Kp)[:, :K]
    W8 = (K + 7) // 8
    qp = np.zeros((N, W8 * 8), np.uint32)
    qp[:, :K] = q
    qweight = np.zeros((N, W8), np.uint32)
    for j in range(8):
        qweight |= qp[:, j::8] << np.uint32(4 * j)
    Z8 = (G + 7) // 8
    zp = np.zeros((N, Z8 * 8), np.uint32)
    zp[:, :G] = zero.astype(np.uint32)
    qzeros = np.zeros((N, Z8), np.uint32)
    for j in range(8):
        qzeros |= zp[:, j::8] << np.uint32(4 * j)
    return qweight, scale16, qzeros


def dequantize_int4_g128(qweight, scales, qzeros, K: int, group: int = 128) -> np.ndarray:
    """w = fp32(scale) * (q - zero) in fp32 (caller casts)."""
    N = qweight.shape[0]
    G = scales.shape[1]
    q = np.zeros((N, qweight.shape[1] * 8), F32)
    for j in range(8):
        q[:, j::8] = ((qweight >> np.uint32(4 * j)) & np.uint32(0xF)).astype(F32)
    z = np.zeros((N, qzeros.shape[1] * 8), F32)
    for j in range(8):
        z[:, j::8] = ((qzeros >> np.uint32(4 * j)) & np.uint32(0xF)).astype(F32)
    q = q[:, :K]
    zk = np.repeat(z[:, :G], group, axis=1)[:, :K]
    sk = np.repeat(scales.astype(F32), group, axis=1)[:, :K]
    return (sk * (q - zk)).astype(F32)


# ---------------------------------------------------------------------------------------------
# int8 (LLM.int8(), bnb.nn.Linear8bitLt(has_fp16_weights=False, threshold=6.0))
# ---------------------------------------------------------------------------------------------
def quantize_int8_rowwise(w: np.ndarray):
    """functional.double_quant row statistics: CB = rint(w * 127 / absmax_row) int8, SCB = absmax_row fp32."""
    w = np.asarray(w, F32)
    scb = np.abs(w).max(axis=1).astype(F32)
    with np.errstate(divide="ignore", invalid="ignore"):
        cb = np.rint(w * (F32(127.0) / scb)[:, None])
    cb = np.nan_to_num(cb, nan=0.0).astype(np.int8)
    return cb, scb


def dequantize_int8_rowwise(cb, scb) -> np.ndarray:
    return (cb.astype(F32) * (np.asarray(scb, F32) / F32(127.0))[:, None]).astype(F32)


def linear_int8(x: np.ndarray, cb, scb, bias=None, threshold: float = 6.0) -> np.ndarray:
    """MatMul8bitLt.forward (inference): fp16 activations; columns with any |x| >= threshold go through
    an fp16 matmul against the dequantised weight columns; the rest is row-wise int8 x int8 -> int32,
    scaled by SCA*SCB/127^2.  x: (M, K) float (treated as fp16), returns (M, N) float32 image of fp16.
    """
    x16 = np.asarray(x, np.float16)
    xf = x16.astype(F32)
    outlier_cols = (np.abs(xf) >= threshold).any(axis=0) if threshold > 0 else np.zeros(xf.shape[1], bool)
    xi = xf.copy()
    xi[:, outlier_cols] = 0
    sca = np.abs(xi).max(axis=1).astype(F32)
    with np.errstate(divide="ignore", invalid="ignore"):
        ca = np.nan_to_num(np.rint(xi * (F32(127.0) / sca)[:, None]), nan=0.0).astype(np.int32)
    out32 = ca @ cb.astype(np.int32).T
    out = out32.astype(F32) * (sca[:, None] * np.asarray(scb, F32)[None, :] / F32(127.0 * 127.0))
    out = _f16(out)
    if bias is not None:
        out = _f16(out.astype(F32) + np.asarray(bias, F32))
    if outlier_cols.any():
        sub_b = _f16(cb[:, outlier_cols].astype(F32) * (np.asarray(scb, F32) / F32(127.0))[:, None])  # (N, n_out)
        sub = xf[:, outlier_cols] @ sub_b.astype(F32).T
        out = _f16(out.astype(F32) + _f16(sub).astype(F32))
    return out.astype(F32)
