"""The numpy and C restatements of the quantisation formats agree bit for bit, and obey the format's own
invariants.  (Parity against bitsandbytes itself is UNPINNED: see oracle/quant_ref.py header.)"""
import numpy as np
import pytest

from oracle import quant_ref as Q
from oracle import quant_ref_c as C


def _w(shape, seed=0, scale=0.02):
    return (np.random.default_rng(seed).standard_normal(shape) * scale).astype(np.float32)


@pytest.mark.parametrize("qt", ["nf4", "fp4"])
@pytest.mark.parametrize("shape", [(64, 256), (7, 33), (1, 64), (3, 1), (128, 4096)])
def test_4bit_numpy_equals_c(qt, shape):
    w = _w(shape, seed=sum(shape))
    p, a = Q.quantize_4bit(w, qt)
    p2, a2 = C.quantize_4bit(w, qt)
    assert p.shape == ((w.size + 1) // 2, 1) and p.dtype == np.uint8  # bnb Params4bit storage shape
    assert np.array_equal(p, p2) and np.array_equal(a, a2)
    d, d2 = Q.dequantize_4bit(p, a, shape, qt), C.dequantize_4bit(p, a, shape, qt)
    assert np.array_equal(d, d2)


@pytest.mark.parametrize("qt", ["nf4", "fp4"])
def test_4bit_format_invariants(qt):
    w = _w((32, 128), 3)
    p, a = Q.quantize_4bit(w, qt)
    # absmax is the block max; even element sits in the HIGH nibble
    assert np.array_equal(a, np.abs(w.reshape(-1, 64)).max(1))
    codes = Q.unpack_codes(p, w.size)
    assert codes[0] == p[0, 0] >> 4 and codes[1] == p[0, 0] & 15
    # every value decodes to the nearest codebook entry (ties aside), extremes are exact
    d = Q.dequantize_4bit(p, a, w.shape, qt)
    code = Q.CODEBOOKS[qt]
    xn = (w.reshape(-1, 64) / a[:, None]).reshape(-1)
    best = np.abs(xn[:, None] - code[None, :]).min(1)
    got = np.abs(xn - code[codes])
    assert np.all(got <= best + 2e-6)  # fp4 thresholds are 6-7 digit decimals in the published tree
    i = np.abs(w).argmax()
    assert d.reshape(-1)[i] == w.reshape(-1)[i]
    # idempotence: quantising the dequantised tensor reproduces the same bytes
    p2, a2 = Q.quantize_4bit(d, qt)
    assert np.array_equal(a, a2)
    if qt == "nf4":
        assert np.array_equal(p, p2)
    else:  # fp4 has a -0 code (0b1000) that re-quantises to +0 (0b0000): equal values, not equal bytes
        assert np.array_equal(Q.dequantize_4bit(p2, a2, w.shape, qt), d)


def test_nf4_known_answers():
    # thresholds are the codebook midpoints; x == threshold goes to the LOWER code (`>` in the tree)
    assert np.allclose((Q.NF4_CODE[:-1] + Q.NF4_CODE[1:]) / 2, Q.NF4_THRESHOLDS, atol=1e-7)
    blk = np.zeros(64, np.float32)
    blk[0], blk[1], blk[2], blk[3] = 1.0, -1.0, Q.NF4_THRESHOLDS[7], np.nextafter(Q.NF4_THRESHOLDS[7], np.float32(1))
    p, a = Q.quantize_4bit(blk)
    assert a[0] == 1.0 and p[0, 0] == 0xF0 and p[1, 0] == 0x78 and p[2, 0] == 0x77
    # all-zero block: 1/0 = inf, 0*inf = NaN, NaN fails every `>` -> code 0, absmax 0 -> dequant -0.0
    p, a = Q.quantize_4bit(np.zeros(64, np.float32))
    assert a[0] == 0 and np.all(p == 0) and np.all(Q.dequantize_4bit(p, a, (64,)) == 0)
    pc, ac = C.quantize_4bit(np.zeros(64, np.float32))
    assert np.array_equal(p, pc) and np.array_equal(a, ac)


def test_dynamic_map_and_double_quant():
    m = Q.DYNAMIC_MAP
    assert m.shape == (256,) and m[-1] == 1.0 and np.all(np.diff(m) > 0) and (m == 0).sum() == 1
    assert np.isclose(m[0], -0.99296875) and np.isclose(m[-2], 0.99296875)
    a = np.abs(_w(1000, 5)) + 0.01
    q, a2, off = Q.double_quantize_absmax(a)
    assert q.dtype == np.uint8 and a2.shape == (4,)
    back = Q.dequantize_absmax(q, a2, off)
    assert np.abs(back - a).max() < 0.02 * a.max()
    # 8-bit code search returns the nearest code
    x = np.linspace(-1, 1, 777).astype(np.float32)
    idx = Q._dquantize8(m, x)
    assert np.all(np.abs(m[idx] - x) <= np.abs(m[None, :] - x[:, None]).min(1) + 1e-7)


@pytest.mark.parametrize("shape", [(16, 256), (5, 200), (8, 128), (3, 8)])
def test_int4_g128(shape):
    w = _w(shape, 11)
    qw, s, z = Q.quantize_int4_g128(w)
    N, K = shape
    assert qw.shape == (N, (K + 7) // 8) and s.shape == (N, (K + 127) // 128) and s.dtype == np.float16
    d = Q.dequantize_int4_g128(qw, s, z, K)
    assert np.array_equal(d, C.dequantize_int4(qw, s, z, K))
    assert np.abs(d - w).max() <= 0.51 * s.astype(np.float32).max() + 1e-8
    qw2, s2, z2 = Q.quantize_int4_g128(d)  # idempotent on its own output up to the scale rounding
    assert np.abs(Q.dequantize_int4_g128(qw2, s2, z2, K) - d).max() <= s.astype(np.float32).max()


def test_gemv_c_matches_numpy():
    w = _w((96, 320), 2)
    x = _w((3, 320), 4, 1.0)
    p, a = Q.quantize_4bit(w)
    y = C.gemv_4bit(p, a, x, 96, 320, round_w_bf16=False)
    assert np.allclose(y, x @ Q.dequantize_4bit(p, a, w.shape).T, rtol=1e-5, atol=1e-6)
    qw, s, z = Q.quantize_int4_g128(w)
    y = C.gemv_int4(qw, s, z, x, 320, round_w_bf16=False)
    assert np.allclose(y, x @ Q.dequantize_int4_g128(qw, s, z, 320).T, rtol=1e-5, atol=1e-6)


def test_int8():
    w = _w((24, 64), 8)
    cb, scb = Q.quantize_int8_rowwise(w)
    assert cb.dtype == np.int8 and np.abs(cb).max() == 127
    assert np.abs(Q.dequantize_int8_rowwise(cb, scb) - w).max() <= scb.max() / 127 * 0.51
    x = _w((2, 64), 9, 1.0)
    x[1, 3] = 7.5  # outlier column -> fp16 side path
    y = Q.linear_int8(x, cb, scb)
    ref = x.astype(np.float16).astype(np.float32) @ w.T
    assert np.abs(y - ref).max() < 0.05
