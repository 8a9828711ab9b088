// On-device weight quantisation / dequantisation.
//   NF4 / FP4 block-wise 4-bit: restates bitsandbytes 0.41 kQuantizeBlockwise / kDequantizeBlockwise (the code the
//   reference reaches through Params4bit.cuda(), generate/base.py:168, generate/tp.py:190); bit-exact to
//   oracle/quant_ref.{py,c}.
//   INT4 group-128: this project's own format, bit-exact to oracle/quant_ref.py::quantize_int4_g128.
// These run once at load time; they are HBM-bound streaming kernels (one warp per quantisation block).
#include "common.cuh"

namespace b200 {

__device__ __forceinline__ unsigned char encode_nf4(float x) {
  // bitsandbytes dQuantizeNF4 decision tree (`>` tests; NaN falls through to code 0)
  if (x > 0.03979014977812767f) {
    if (x > 0.3893125355243683f) {
      if (x > 0.6427869200706482f) return x > 0.8614784181118011f ? 15 : 14;
      return x > 0.5016634166240692f ? 13 : 12;
    }
    if (x > 0.2035212516784668f) return x > 0.2920137718319893f ? 11 : 10;
    return x > 0.1202552504837513f ? 9 : 8;
  }
  if (x > -0.33967943489551544f) {
    if (x > -0.13791173323988914f) return x > -0.045525018125772476f ? 7 : 6;
    return x > -0.23460740596055984f ? 5 : 4;
  }
  if (x > -0.6106329262256622f) return x > -0.4599952697753906f ? 3 : 2;
  return x > -0.8480964004993439f ? 1 : 0;
}

__device__ __forceinline__ unsigned char encode_fp4(float x) {
  // bitsandbytes dQuantizeFP4
  int sign = x < 0.f ? 8 : 0;
  x = fabsf(x);
  if (x > 0.29166667f) {
    if (x > 0.583333f) return (x > 0.8333333f ? 3 : 2) + sign;
    return (x > 0.4166667f ? 5 : 4) + sign;
  }
  if (x > 0.0859375f) return (x > 0.20833333f ? 7 : 6) + sign;
  return (x > 0.00260417f ? 1 : 0) + sign;
}

// one warp per quantisation block; lane handles element pairs (2*lane + 64*j, +1) -> one packed byte each
template <typename T>
__global__ void __launch_bounds__(256) quantize_4bit_kernel(const T* __restrict__ w, int64_t numel, int blocksize, int qtype,
                                                            uint8_t* __restrict__ packed, float* __restrict__ absmax) {
  griddep_wait();
  const int lane = threadIdx.x & 31;
  const int64_t nblocks = (numel + blocksize - 1) / blocksize;
  const int64_t warps_total = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t blk = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); blk < nblocks; blk += warps_total) {
    const int64_t base = blk * blocksize;
    float am = 0.f;
    for (int i = 2 * lane; i < blocksize; i += 64) {
      int64_t e = base + i;
      if (e < numel) am = fmaxf(am, fabsf(load_as_float(w, e)));
      if (e + 1 < numel) am = fmaxf(am, fabsf(load_as_float(w, e + 1)));
    }
    am = warp_max(am);
    if (lane == 0) absmax[blk] = am;
    const float inv = 1.0f / am;  // reciprocal then multiply, as the reference kernel does
    for (int i = 2 * lane; i < blocksize; i += 64) {
      int64_t e = base + i;
      if (e >= numel) break;
      float a = load_as_float(w, e) * inv;
      unsigned char hi = qtype == B200_Q_NF4 ? encode_nf4(a) : encode_fp4(a);
      unsigned char lo = 0;
      if (e + 1 < numel) {
        float b = load_as_float(w, e + 1) * inv;
        lo = qtype == B200_Q_NF4 ? encode_nf4(b) : encode_fp4(b);
      }
      packed[e >> 1] = (unsigned char)((hi << 4) | lo);
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256) dequantize_4bit_kernel(const uint8_t* __restrict__ packed, const float* __restrict__ absmax,
                                                              int64_t numel, int blocksize, int qtype, T* __restrict__ out) {
  griddep_wait();
  const int64_t nbytes = (numel + 1) / 2;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nbytes; i += stride) {
    unsigned char b = packed[i];
    int64_t e = 2 * i;
    float s = absmax[e / blocksize];  // blocksize even: both nibbles share a block
    store_from_float(out, e, code_value(qtype, b >> 4) * s);
    if (e + 1 < numel) store_from_float(out, e + 1, code_value(qtype, b & 15) * s);
  }
}

// ---- int4 group quantisation: one warp per (row, 8 consecutive groups) so it owns one qzeros word -----------
template <typename T>
__global__ void __launch_bounds__(256) quantize_int4_kernel(const T* __restrict__ w, int64_t N, int64_t K, int group,
                                                            uint32_t* __restrict__ qweight, uint16_t* __restrict__ scales,
                                                            uint32_t* __restrict__ qzeros, uint32_t* __restrict__ sz) {
  griddep_wait();
  const int lane = threadIdx.x & 31;
  const int64_t G = (K + group - 1) / group, Z8 = (G + 7) / 8, W8 = (K + 7) / 8;
  const int64_t items = N * Z8;
  const int64_t warps_total = (int64_t)gridDim.x * (blockDim.x >> 5);
  const int per_lane = group / 32;  // group is a multiple of 32 with per_lane in {4, 8}; host checks
  for (int64_t it = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); it < items; it += warps_total) {
    const int64_t n = it / Z8, zw = it % Z8;
    uint32_t zword = 0;
    for (int gi = 0; gi < 8; ++gi) {
      const int64_t g = zw * 8 + gi;
      if (g >= G) break;
      const int64_t k0 = g * group + (int64_t)lane * per_lane;
      float v[8];
      float lo = 0.f, hi = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        v[j] = 0.f;
        if (j < per_lane && k0 + j < K) v[j] = load_as_float(w, n * K + k0 + j);
        lo = fminf(lo, v[j]);
        hi = fmaxf(hi, v[j]);
      }
      lo = -warp_max(-lo);
      hi = warp_max(hi);
      float scale = (hi - lo) / 15.0f;
      if (scale == 0.f) scale = 1.0f;
      __half s16 = __float2half_rn(scale);
      if (__half2float(s16) == 0.f) s16 = __ushort_as_half((unsigned short)0x0400);  // 2^-14 underflow guard
      const float s = __half2float(s16);
      const float zero = fminf(fmaxf(rintf((-lo) / s), 0.f), 15.f);
      uint32_t nib = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (j < per_lane) {
          float q = fminf(fmaxf(rintf(v[j] / s) + zero, 0.f), 15.f);
          if (k0 + j >= K) q = 0.f;
          nib |= ((uint32_t)q) << (4 * j);
        }
      }
      if (per_lane == 4) {  // two lanes form one word: even lane low half, odd lane high half
        uint32_t other = __shfl_xor_sync(0xffffffffu, nib, 1);
        uint32_t word = (lane & 1) ? ((nib << 16) | (other & 0xffffu)) : ((other << 16) | (nib & 0xffffu));
        int64_t widx = (g * group) / 8 + (lane >> 1);
        if (!(lane & 1) && widx < W8) qweight[n * W8 + widx] = word;
      } else {
        int64_t widx = (g * group) / 8 + lane;
        if (widx < W8) qweight[n * W8 + widx] = nib;
      }
      if (lane == 0) {
        scales[n * G + g] = __half_as_ushort(s16);
        if (sz) sz[n * G + g] = (uint32_t)__half_as_ushort(s16) | ((uint32_t)__half_as_ushort(__float2half_rn(zero)) << 16);
      }
      zword |= ((uint32_t)zero) << (4 * gi);
    }
    if (lane == 0) qzeros[n * Z8 + zw] = zword;
  }
}

template <typename T>
__global__ void __launch_bounds__(256) dequantize_int4_kernel(const uint32_t* __restrict__ qweight, const uint16_t* __restrict__ scales,
                                                              const uint32_t* __restrict__ qzeros, int64_t N, int64_t K, int group,
                                                              T* __restrict__ out) {
  griddep_wait();
  const int64_t G = (K + group - 1) / group, Z8 = (G + 7) / 8, W8 = (K + 7) / 8;
  const int64_t total = N * K;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int64_t n = i / K, k = i % K, g = k / group;
    float q = (float)((qweight[n * W8 + k / 8] >> (4 * (k % 8))) & 0xF);
    float z = (float)((qzeros[n * Z8 + g / 8] >> (4 * (g % 8))) & 0xF);
    float s = __half2float(__ushort_as_half(scales[n * G + g]));
    store_from_float(out, i, s * (q - z));
  }
}


// ---- double quantisation of the absmax vector (bnb.nf4-dq / fp4-dq): per 256 values, subtract the
// caller-given offset (mean), take the block absmax, map x / absmax2 to the nearest entry of the 256-entry
// "dynamic" code with bitsandbytes' pivot search (kernels.cu dQuantize<0>), and write back the DEQUANTISED value
// code[q] * absmax2 + offset, which is what every later dequantise / GEMV of a -dq layer uses.
__global__ void __launch_bounds__(256) double_quant_absmax_kernel(const float* __restrict__ absmax, int64_t n, const float* __restrict__ code,
                                                                  float offset, uint8_t* __restrict__ q_out, float* __restrict__ absmax2_out,
                                                                  float* __restrict__ deq_out) {
  __shared__ float s_code[256];
  __shared__ float s_max[8];
  griddep_wait();
  const int tid = threadIdx.x;
  s_code[tid] = code[tid];
  const int64_t i = (int64_t)blockIdx.x * 256 + tid;
  const float v = i < n ? __fsub_rn(absmax[i], offset) : 0.f;
  float m = warp_max(fabsf(v));
  if ((tid & 31) == 0) s_max[tid >> 5] = m;
  __syncthreads();
  m = s_max[0];
#pragma unroll
  for (int w = 1; w < 8; ++w) m = fmaxf(m, s_max[w]);
  if (tid == 0) absmax2_out[blockIdx.x] = m;
  if (i >= n) return;
  const float x = __fmul_rn(v, __fdiv_rn(1.0f, m));
  int pivot = 127, upper_pivot = 255, lower_pivot = 0;
  float lower = -1.0f, upper = 1.0f, val = s_code[pivot];
  for (int st = 64; st > 0; st >>= 1) {
    if (x > val) { lower_pivot = pivot; lower = val; pivot += st; }
    else { upper_pivot = pivot; upper = val; pivot -= st; }
    val = s_code[pivot];
  }
  if (upper_pivot == 255) upper = s_code[upper_pivot];
  if (lower_pivot == 0) lower = s_code[lower_pivot];
  int q;
  if (x > val) q = x > __fmul_rn(__fadd_rn(upper, val), 0.5f) ? upper_pivot : pivot;
  else q = x < __fmul_rn(__fadd_rn(lower, val), 0.5f) ? lower_pivot : pivot;
  q_out[i] = (uint8_t)q;
  deq_out[i] = __fadd_rn(__fmul_rn(s_code[q], m), offset);
}

static int grid_for(int64_t work_items, int per_block) {
  int64_t blocks = (work_items + per_block - 1) / per_block;
  int64_t cap = (int64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

}  // namespace b200

using namespace b200;

extern "C" {

int b200_quantize_4bit(const void* w, int w_dtype, int64_t numel, int blocksize, int qtype, uint8_t* packed, float* absmax,
                       b200_stream_t stream) {
  B200_CHECK_ARG(w && packed && absmax, "b200_quantize_4bit: null pointer");
  B200_CHECK_ARG(numel > 0, "b200_quantize_4bit: numel must be positive (got %lld)", (long long)numel);
  B200_CHECK_ARG(blocksize >= 2 && blocksize % 2 == 0, "b200_quantize_4bit: blocksize must be even (got %d)", blocksize);
  B200_CHECK_ARG(qtype == B200_Q_NF4 || qtype == B200_Q_FP4, "b200_quantize_4bit: qtype must be NF4 or FP4");
  cudaStream_t st = (cudaStream_t)stream;
  int64_t nblocks = (numel + blocksize - 1) / blocksize;
  int grid = grid_for(nblocks, 8);
  switch (w_dtype) {
    case B200_BF16: B200_CUDA(launch(quantize_4bit_kernel<__nv_bfloat16>, dim3(grid), dim3(256), 0, st, (const __nv_bfloat16*)w, numel, blocksize, qtype, packed, absmax)); break;
    case B200_F16: B200_CUDA(launch(quantize_4bit_kernel<__half>, dim3(grid), dim3(256), 0, st, (const __half*)w, numel, blocksize, qtype, packed, absmax)); break;
    case B200_F32: B200_CUDA(launch(quantize_4bit_kernel<float>, dim3(grid), dim3(256), 0, st, (const float*)w, numel, blocksize, qtype, packed, absmax)); break;
    default: B200_CHECK_ARG(false, "b200_quantize_4bit: unsupported dtype %d", w_dtype);
  }
  return B200_OK;
}

int b200_dequantize_4bit(const uint8_t* packed, const float* absmax, int64_t numel, int blocksize, int qtype, void* out,
                         int out_dtype, b200_stream_t stream) {
  B200_CHECK_ARG(out && packed && absmax, "b200_dequantize_4bit: null pointer");
  B200_CHECK_ARG(numel > 0, "b200_dequantize_4bit: numel must be positive");
  B200_CHECK_ARG(blocksize >= 2 && blocksize % 2 == 0, "b200_dequantize_4bit: blocksize must be even (got %d)", blocksize);
  B200_CHECK_ARG(qtype == B200_Q_NF4 || qtype == B200_Q_FP4, "b200_dequantize_4bit: qtype must be NF4 or FP4");
  cudaStream_t st = (cudaStream_t)stream;
  int grid = grid_for((numel + 1) / 2, 256 * 4);
  switch (out_dtype) {
    case B200_BF16: B200_CUDA(launch(dequantize_4bit_kernel<__nv_bfloat16>, dim3(grid), dim3(256), 0, st, packed, absmax, numel, blocksize, qtype, (__nv_bfloat16*)out)); break;
    case B200_F16: B200_CUDA(launch(dequantize_4bit_kernel<__half>, dim3(grid), dim3(256), 0, st, packed, absmax, numel, blocksize, qtype, (__half*)out)); break;
    case B200_F32: B200_CUDA(launch(dequantize_4bit_kernel<float>, dim3(grid), dim3(256), 0, st, packed, absmax, numel, blocksize, qtype, (float*)out)); break;
    default: B200_CHECK_ARG(false, "b200_dequantize_4bit: unsupported dtype %d", out_dtype);
  }
  return B200_OK;
}

int b200_quantize_int4(const void* w, int w_dtype, int64_t N, int64_t K, int group, uint32_t* qweight, uint16_t* scales,
                       uint32_t* qzeros, uint32_t* sz, b200_stream_t stream) {
  B200_CHECK_ARG(w && qweight && scales && qzeros, "b200_quantize_int4: null pointer");
  B200_CHECK_ARG(N > 0 && K > 0, "b200_quantize_int4: empty weight");
  B200_CHECK_ARG(group == 128 || group == 256, "b200_quantize_int4: group must be 128 or 256 (got %d)", group);
  cudaStream_t st = (cudaStream_t)stream;
  int64_t G = (K + group - 1) / group, Z8 = (G + 7) / 8;
  int grid = grid_for(N * Z8, 8);
  switch (w_dtype) {
    case B200_BF16: B200_CUDA(launch(quantize_int4_kernel<__nv_bfloat16>, dim3(grid), dim3(256), 0, st, (const __nv_bfloat16*)w, N, K, group, qweight, scales, qzeros, sz)); break;
    case B200_F16: B200_CUDA(launch(quantize_int4_kernel<__half>, dim3(grid), dim3(256), 0, st, (const __half*)w, N, K, group, qweight, scales, qzeros, sz)); break;
    case B200_F32: B200_CUDA(launch(quantize_int4_kernel<float>, dim3(grid), dim3(256), 0, st, (const float*)w, N, K, group, qweight, scales, qzeros, sz)); break;
    default: B200_CHECK_ARG(false, "b200_quantize_int4: unsupported dtype %d", w_dtype);
  }
  return B200_OK;
}

int b200_dequantize_int4(const uint32_t* qweight, const uint16_t* scales, const uint32_t* qzeros, int64_t N, int64_t K,
                         int group, void* out, int out_dtype, b200_stream_t stream) {
  B200_CHECK_ARG(out && qweight && scales && qzeros, "b200_dequantize_int4: null pointer");
  B200_CHECK_ARG(N > 0 && K > 0 && group > 0, "b200_dequantize_int4: bad shape");
  cudaStream_t st = (cudaStream_t)stream;
  int grid = grid_for(N * K, 256 * 4);
  switch (out_dtype) {
    case B200_BF16: B200_CUDA(launch(dequantize_int4_kernel<__nv_bfloat16>, dim3(grid), dim3(256), 0, st, qweight, scales, qzeros, N, K, group, (__nv_bfloat16*)out)); break;
    case B200_F16: B200_CUDA(launch(dequantize_int4_kernel<__half>, dim3(grid), dim3(256), 0, st, qweight, scales, qzeros, N, K, group, (__half*)out)); break;
    case B200_F32: B200_CUDA(launch(dequantize_int4_kernel<float>, dim3(grid), dim3(256), 0, st, qweight, scales, qzeros, N, K, group, (float*)out)); break;
    default: B200_CHECK_ARG(false, "b200_dequantize_int4: unsupported dtype %d", out_dtype);
  }
  return B200_OK;
}

int b200_double_quant_absmax(const float* absmax, int64_t n, const float* code256, float offset, uint8_t* q_out, float* absmax2_out,
                             float* absmax_deq_out, b200_stream_t stream) {
  B200_CHECK_ARG(absmax && code256 && q_out && absmax2_out && absmax_deq_out, "b200_double_quant_absmax: null pointer");
  B200_CHECK_ARG(n > 0 && (n + 255) / 256 <= 0x7fffffff, "b200_double_quant_absmax: bad length %lld", (long long)n);
  B200_CUDA(launch(double_quant_absmax_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, (cudaStream_t)stream, absmax, n, code256, offset,
                   q_out, absmax2_out, absmax_deq_out));
  return B200_OK;
}

}  // extern "C"
