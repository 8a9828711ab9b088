/* CPU restatement (plain C, OpenMP) of the quantised-Linear arithmetic on the hot path.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): used by tests/ as the checker and by bench.py as the
 * timed CPU baseline ("port").  The product library never links or loads this file.
 *
 * PARITY UNPINNED: the arithmetic restated here lives in bitsandbytes==0.41.0 (reference
 * requirements-all.txt:3; call sites generate/base.py:129-134, generate/tp.py:127-134,
 * lit_gpt/lora.py:156-160), which is absent from /root/reference.  It follows the published kernels of
 * that version: kQuantizeBlockwise / kDequantizeBlockwise / dQuantizeNF4 / dQuantizeFP4 /
 * dDequantizeFP4Tree (csrc/kernels.cu) and kgemm_4bit_inference_naive for the GEMV.  int4 group-128 is
 * this project's own format.  Cross-checked against oracle/quant_ref.py in tests/test_oracle_quant.py.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static const float NF4_CODE[16] = {
    -1.0f, -0.6961928009986877f, -0.5250730514526367f, -0.39491748809814453f, -0.28444138169288635f,
    -0.18477343022823334f, -0.09105003625154495f, 0.0f, 0.07958029955625534f, 0.16093020141124725f,
    0.24611230194568634f, 0.33791524171829224f, 0.44070982933044434f, 0.5626170039176941f,
    0.7229568362236023f, 1.0f};

static const float FP4_CODE[16] = {
    0.0f, 5.208333333e-03f, 0.66666667f, 1.0f, 0.33333333f, 0.5f, 0.16666667f, 0.25f,
    -0.0f, -5.208333333e-03f, -0.66666667f, -1.0f, -0.33333333f, -0.5f, -0.16666667f, -0.25f};

/* kernels.cu dQuantizeNF4 */
static unsigned char q_nf4(float x) {
  if (x > 0.03979014977812767f)
    if (x > 0.3893125355243683f)
      if (x > 0.6427869200706482f)
        return x > 0.8614784181118011f ? 15 : 14;
      else
        return x > 0.5016634166240692f ? 13 : 12;
    else if (x > 0.2035212516784668f)
      return x > 0.2920137718319893f ? 11 : 10;
    else
      return x > 0.1202552504837513f ? 9 : 8;
  else if (x > -0.33967943489551544f)
    if (x > -0.13791173323988914f)
      return x > -0.045525018125772476f ? 7 : 6;
    else
      return x > -0.23460740596055984f ? 5 : 4;
  else if (x > -0.6106329262256622f)
    return x > -0.4599952697753906f ? 3 : 2;
  else
    return x > -0.8480964004993439f ? 1 : 0;
}

/* kernels.cu dQuantizeFP4 */
static unsigned char q_fp4(float x) {
  int sign = x < 0 ? 8 : 0;
  x = fabsf(x);
  if (x > 0.29166667f)
    if (x > 0.583333f)
      return (x > 0.8333333f ? 3 : 2) + sign;
    else
      return (x > 0.4166667f ? 5 : 4) + sign;
  else if (x > 0.0859375f)
    return (x > 0.20833333f ? 7 : 6) + sign;
  else
    return (x > 0.00260417f ? 1 : 0) + sign;
}

/* quant_type: 0 = nf4, 1 = fp4 */
void ref_quantize_4bit(const float* w, int64_t n, int blocksize, int quant_type, uint8_t* packed, float* absmax) {
  int64_t nblocks = (n + blocksize - 1) / blocksize;
  memset(packed, 0, (size_t)((n + 1) / 2));
#pragma omp parallel for schedule(static)
  for (int64_t b = 0; b < nblocks; ++b) {
    int64_t lo = b * blocksize, hi = lo + blocksize < n ? lo + blocksize : n;
    float am = 0.f;
    for (int64_t i = lo; i < hi; ++i) am = fmaxf(am, fabsf(w[i]));
    absmax[b] = am;
    float inv = 1.0f / am;
    for (int64_t i = lo; i < hi; ++i) {
      float xn = w[i] * inv;
      unsigned char c = quant_type == 0 ? q_nf4(xn) : q_fp4(xn);
      /* blocksize is even, so a byte never straddles two blocks: no race */
      packed[i >> 1] |= (i & 1) ? c : (unsigned char)(c << 4);
    }
  }
}

void ref_dequantize_4bit(const uint8_t* packed, const float* absmax, int64_t n, int blocksize, int quant_type, float* out) {
  const float* code = quant_type == 0 ? NF4_CODE : FP4_CODE;
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) {
    unsigned char byte = packed[i >> 1];
    unsigned char c = (i & 1) ? (byte & 0x0F) : (byte >> 4);
    out[i] = code[c] * absmax[i / blocksize];
  }
}

/* round-to-nearest-even float -> bf16 -> float (the compute dtype of the reference's bf16-true path) */
static inline float bf16_round(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return f; /* NaN */
  u += 0x7fffu + ((u >> 16) & 1u);
  u &= 0xffff0000u;
  memcpy(&f, &u, 4);
  return f;
}

/* y[b, n] = sum_k x[b, k] * bf16(code[q[n,k]] * absmax[(n*K+k)/bs])   (fp32 accumulate; y left in fp32).
 * This is "dequantize_4bit to the compute dtype, then F.linear" (the reference's batch>1 / prefill path,
 * SURVEY Q1) restated with a wide accumulator; it is also the timed CPU baseline for the GEMV. */
void ref_gemv_4bit(const uint8_t* packed, const float* absmax, const float* x, float* y, int64_t N, int64_t K, int B,
                   int blocksize, int quant_type, int round_w_bf16) {
  const float* code = quant_type == 0 ? NF4_CODE : FP4_CODE;
#pragma omp parallel for schedule(static)
  for (int64_t n = 0; n < N; ++n) {
    double acc[64];
    for (int b = 0; b < B; ++b) acc[b] = 0.0;
    for (int64_t k = 0; k < K; ++k) {
      int64_t i = n * K + k;
      unsigned char byte = packed[i >> 1];
      unsigned char c = (i & 1) ? (byte & 0x0F) : (byte >> 4);
      float wv = code[c] * absmax[i / blocksize];
      if (round_w_bf16) wv = bf16_round(wv);
      for (int b = 0; b < B; ++b) acc[b] += (double)wv * (double)x[(int64_t)b * K + k];
    }
    for (int b = 0; b < B; ++b) y[(int64_t)b * N + n] = (float)acc[b];
  }
}

/* int4 group-128 (project format, see oracle/quant_ref.py::quantize_int4_g128): dequant + GEMV */
static inline float half_to_float(uint16_t h) {
  uint32_t s = (uint32_t)(h & 0x8000u) << 16, e = (h >> 10) & 0x1f, m = h & 0x3ff, u;
  if (e == 0) {
    if (m == 0) u = s;
    else {
      int sh = 0;
      while (!(m & 0x400)) { m <<= 1; ++sh; }
      u = s | ((uint32_t)(113 - sh) << 23) | ((m & 0x3ff) << 13);
    }
  } else if (e == 31) u = s | 0x7f800000u | (m << 13);
  else u = s | ((e + 112) << 23) | (m << 13);
  float f;
  memcpy(&f, &u, 4);
  return f;
}

void ref_dequantize_int4(const uint32_t* qweight, const uint16_t* scales, const uint32_t* qzeros, int64_t N, int64_t K,
                         int group, float* out) {
  int64_t W8 = (K + 7) / 8, G = (K + group - 1) / group, Z8 = (G + 7) / 8;
#pragma omp parallel for schedule(static)
  for (int64_t n = 0; n < N; ++n)
    for (int64_t k = 0; k < K; ++k) {
      int64_t g = k / group;
      float q = (float)((qweight[n * W8 + k / 8] >> (4 * (k % 8))) & 0xF);
      float z = (float)((qzeros[n * Z8 + g / 8] >> (4 * (g % 8))) & 0xF);
      out[n * K + k] = half_to_float(scales[n * G + g]) * (q - z);
    }
}

void ref_gemv_int4(const uint32_t* qweight, const uint16_t* scales, const uint32_t* qzeros, const float* x, float* y,
                   int64_t N, int64_t K, int B, int group, int round_w_bf16) {
  int64_t W8 = (K + 7) / 8, G = (K + group - 1) / group, Z8 = (G + 7) / 8;
#pragma omp parallel for schedule(static)
  for (int64_t n = 0; n < N; ++n) {
    double acc[64];
    for (int b = 0; b < B; ++b) acc[b] = 0.0;
    for (int64_t k = 0; k < K; ++k) {
      int64_t g = k / group;
      float q = (float)((qweight[n * W8 + k / 8] >> (4 * (k % 8))) & 0xF);
      float z = (float)((qzeros[n * Z8 + g / 8] >> (4 * (g % 8))) & 0xF);
      float wv = half_to_float(scales[n * G + g]) * (q - z);
      if (round_w_bf16) wv = bf16_round(wv);
      for (int b = 0; b < B; ++b) acc[b] += (double)wv * (double)x[(int64_t)b * K + k];
    }
    for (int b = 0; b < B; ++b) y[(int64_t)b * N + n] = (float)acc[b];
  }
}

int ref_num_threads(void) {
#ifdef _OPENMP
  extern int omp_get_max_threads(void);
  return omp_get_max_threads();
#else
  return 1;
#endif
}
