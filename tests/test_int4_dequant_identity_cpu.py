"""The prefill GEMM's int4 dequantisation (csrc/qgemm.cu) replaces  scale * (float(q) - zero)  by ONE fused multiply-add
fma(scale, 16 + q, -scale * (16 + zero))  with  16 + q  assembled from the nibble by bit operations.  This checks, for EVERY fp16 scale and
every (q, zero) pair, that the two are the same fp32 number (so the bf16 weight the tensor core sees is unchanged), and that the bit pattern
used for 16 + q is right.  Pure numpy: no GPU, no library."""
import numpy as np


def test_nibble_in_mantissa_is_16_plus_q():
    q = np.arange(16, dtype=np.uint32)
    bits = (q << np.uint32(19)) | np.uint32(0x41800000)          # nibble in the top four mantissa bits of 16.0f
    assert np.array_equal(bits.view(np.float32), (16 + q).astype(np.float32))


def test_fma_form_equals_reference_form_for_every_fp16_scale():
    scale16 = np.arange(65536, dtype=np.uint16).view(np.float16)
    scale16 = scale16[np.isfinite(scale16)]
    s32 = scale16.astype(np.float32)
    s64 = s32.astype(np.float64)
    for zero in range(17):                                       # stored zero points are integers in [0, 16]
        c0 = (-(s32 * np.float32(16 + zero))).astype(np.float32)  # exact: 11-bit significand x 6-bit integer
        assert np.array_equal(c0.astype(np.float64), -(s64 * (16 + zero)))
        for q in range(16):
            ref = (s32 * np.float32(q - zero)).astype(np.float32)              # dequantise kernel: fp32 product of fp16 scale and integer
            fma = (s64 * (16 + q) + c0.astype(np.float64)).astype(np.float32)  # FFMA: exact product and sum, one rounding
            assert np.array_equal(ref.view(np.uint32) & 0x7FFFFFFF, fma.view(np.uint32) & 0x7FFFFFFF)  # (+0 / -0 round to the same bf16 magnitude)
            assert np.array_equal(ref, fma)
