// Small latency-bound kernels of the decode step: RMSNorm (+residual add), embedding gather, greedy argmax.
// They exist so that a whole decode step is a CUDA graph made only of this library's kernels, chained with
// programmatic dependent launch.  Bytes are negligible (<= a few hundred KB); launch latency is what counts.
#include "common.cuh"

namespace b200 {

// y = weight * ((x+delta) * rsqrt(mean((x+delta)^2) + eps)), fp32 math, one cast (lit_gpt/rmsnorm.py:19-25).
// One CTA per row; the row is cached in registers between the two passes when it fits (dim <= 8 * 8 * 256).
__global__ void __launch_bounds__(256) rmsnorm_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ delta,
                                                      const __nv_bfloat16* __restrict__ weight, float eps,
                                                      __nv_bfloat16* __restrict__ y, __nv_bfloat16* __restrict__ sum_out,
                                                      int64_t dim) {
  __shared__ float s_part[8];
  __shared__ float s_rstd;
  griddep_launch();
  griddep_wait();
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t row = blockIdx.x;
  const __nv_bfloat16* xr = x + row * dim;
  const __nv_bfloat16* dr = delta ? delta + row * dim : nullptr;
  float ss = 0.f;
  for (int64_t i = tid; i < dim; i += 256) {
    float v = __bfloat162float(xr[i]);
    if (dr) v = round_bf16(v + __bfloat162float(dr[i]));
    if (sum_out) sum_out[row * dim + i] = __float2bfloat16_rn(v);
    ss = fmaf(v, v, ss);
  }
  ss = warp_sum(ss);
  if (lane == 0) s_part[warp] = ss;
  __syncthreads();
  if (tid == 0) {
    float tot = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) tot += s_part[w];
    s_rstd = rsqrtf(tot / (float)dim + eps);
  }
  __syncthreads();
  const float rstd = s_rstd;
  for (int64_t i = tid; i < dim; i += 256) {
    float v = __bfloat162float(xr[i]);
    if (dr) v = round_bf16(v + __bfloat162float(dr[i]));
    y[row * dim + i] = __float2bfloat16_rn(__bfloat162float(weight[i]) * (v * rstd));
  }
}

// y = silu(a) * b elementwise over bf16 (lit_gpt/model.py:715 at prefill; the decode GEMV has it as its epilogue): the reference's
// rounding points -- F.silu rounds to bf16, the product rounds to bf16 -- eight elements per thread and step
__global__ void __launch_bounds__(256) swiglu_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ b,
                                                     __nv_bfloat16* __restrict__ y, int64_t n8) {
  griddep_launch();
  griddep_wait();
  for (int64_t i = blockIdx.x * 256 + threadIdx.x; i < n8; i += (int64_t)gridDim.x * 256) {
    const uint4 av = reinterpret_cast<const uint4*>(a)[i], bv = reinterpret_cast<const uint4*>(b)[i];
    const uint32_t aw[4] = {av.x, av.y, av.z, av.w}, bw[4] = {bv.x, bv.y, bv.z, bv.w};
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float a0 = bf16lo(aw[j]), a1 = bf16hi(aw[j]);
      const float s0 = round_bf16(a0 / (1.0f + expf(-a0))), s1 = round_bf16(a1 / (1.0f + expf(-a1)));
      o[j] = pack_bf16(s0 * bf16lo(bw[j]), s1 * bf16hi(bw[j]));
    }
    reinterpret_cast<uint4*>(y)[i] = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

template <typename IdxT>
__global__ void __launch_bounds__(256) embedding_kernel(const IdxT* __restrict__ idx, const __nv_bfloat16* __restrict__ wte,
                                                        __nv_bfloat16* __restrict__ out, int64_t dim, int64_t vocab) {
  griddep_launch();
  griddep_wait();
  const int64_t row = blockIdx.x;
  int64_t tok = (int64_t)idx[row];
  tok = tok < 0 ? 0 : (tok >= vocab ? vocab - 1 : tok);  // never read out of bounds
  const __nv_bfloat16* src = wte + tok * dim;
  if ((dim & 7) == 0) {
    const uint4* s4 = reinterpret_cast<const uint4*>(src);
    uint4* d4 = reinterpret_cast<uint4*>(out + row * dim);
    for (int64_t i = threadIdx.x; i < (dim >> 3); i += 256) d4[i] = s4[i];
  } else {
    for (int64_t i = threadIdx.x; i < dim; i += 256) out[row * dim + i] = src[i];
  }
}

// one CTA (1024 threads) per batch row; ties resolve to the LOWEST index, as torch.argmax does
template <typename TokT, typename PosT>
__global__ void __launch_bounds__(1024) argmax_kernel(const __nv_bfloat16* __restrict__ logits, int64_t vocab, int64_t stride,
                                                      TokT* __restrict__ token_out, PosT* __restrict__ pos_inout,
                                                      TokT* __restrict__ history, int64_t history_len) {
  __shared__ float s_v[32];
  __shared__ int s_i[32];
  griddep_launch();
  griddep_wait();
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const __nv_bfloat16* lr = logits + (int64_t)blockIdx.x * stride;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  for (int64_t i = tid; i < vocab; i += 1024) {
    float v = __bfloat162float(lr[i]);
    if (v > best || (v == best && (int)i < bi) || bi == 0x7fffffff) { best = v; bi = (int)i; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float ov = __shfl_xor_sync(0xffffffffu, best, o);
    int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
  }
  if (lane == 0) { s_v[warp] = best; s_i[warp] = bi; }
  __syncthreads();
  if (warp == 0) {
    best = s_v[lane];
    bi = s_i[lane];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      float ov = __shfl_xor_sync(0xffffffffu, best, o);
      int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    if (lane == 0) {
      token_out[blockIdx.x] = (TokT)bi;
      if (pos_inout && blockIdx.x == 0) {
        const int64_t np = (int64_t)pos_inout[0] + 1;
        pos_inout[0] = (PosT)np;
        if (history && np >= 0 && np < history_len) history[np] = (TokT)bi;  // sequence index of the sampled token
      }
    }
  }
}

}  // namespace b200

using namespace b200;

extern "C" {

int b200_rmsnorm(const void* x, const void* delta, const void* weight, float eps, void* y, void* sum_out, int64_t rows,
                 int64_t dim, b200_stream_t stream) {
  B200_CHECK_ARG(x && weight && y, "b200_rmsnorm: null pointer");
  B200_CHECK_ARG(rows > 0 && dim > 0, "b200_rmsnorm: empty input (rows=%lld dim=%lld)", (long long)rows, (long long)dim);
  B200_CHECK_ARG(rows <= 0x7fffffff, "b200_rmsnorm: too many rows");
  B200_CUDA(launch(rmsnorm_kernel, dim3((unsigned)rows), dim3(256), 0, (cudaStream_t)stream, (const __nv_bfloat16*)x,
                   (const __nv_bfloat16*)delta, (const __nv_bfloat16*)weight, eps, (__nv_bfloat16*)y, (__nv_bfloat16*)sum_out, dim));
  return B200_OK;
}

int b200_swiglu(const void* a, const void* b, void* y, int64_t n, b200_stream_t stream) {
  B200_CHECK_ARG(a && b && y && n > 0 && n % 8 == 0, "b200_swiglu: null pointer or element count not a multiple of 8");
  B200_CHECK_ARG(((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(y)) & 15) == 0, "b200_swiglu: pointers must be 16-byte aligned");
  const int64_t n8 = n / 8;
  int64_t blocks = (n8 + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  B200_CUDA(launch(swiglu_kernel, dim3((unsigned)blocks), dim3(256), 0, (cudaStream_t)stream, (const __nv_bfloat16*)a, (const __nv_bfloat16*)b, (__nv_bfloat16*)y, n8));
  return B200_OK;
}

int b200_embedding(const void* idx, int idx_is_int64, const void* wte, void* out, int64_t rows, int64_t dim, int64_t vocab,
                   b200_stream_t stream) {
  B200_CHECK_ARG(idx && wte && out, "b200_embedding: null pointer");
  B200_CHECK_ARG(rows > 0 && dim > 0 && vocab > 0 && rows <= 0x7fffffff, "b200_embedding: bad shape");
  cudaStream_t st = (cudaStream_t)stream;
  if (idx_is_int64)
    B200_CUDA(launch(embedding_kernel<int64_t>, dim3((unsigned)rows), dim3(256), 0, st, (const int64_t*)idx, (const __nv_bfloat16*)wte, (__nv_bfloat16*)out, dim, vocab));
  else
    B200_CUDA(launch(embedding_kernel<int32_t>, dim3((unsigned)rows), dim3(256), 0, st, (const int32_t*)idx, (const __nv_bfloat16*)wte, (__nv_bfloat16*)out, dim, vocab));
  return B200_OK;
}

int b200_argmax(const void* logits, int64_t batch, int64_t vocab, int64_t stride, void* token_out, int token_is_int64,
                void* pos_inout, int pos_is_int64, void* history, int64_t history_len, b200_stream_t stream) {
  B200_CHECK_ARG(logits && token_out, "b200_argmax: null pointer");
  B200_CHECK_ARG(batch > 0 && vocab > 0 && stride >= vocab && vocab < 0x7fffffff, "b200_argmax: bad shape");
  cudaStream_t st = (cudaStream_t)stream;
  const __nv_bfloat16* l = (const __nv_bfloat16*)logits;
  dim3 grid((unsigned)batch), block(1024);
  if (token_is_int64 && pos_is_int64) B200_CUDA(launch(argmax_kernel<int64_t, int64_t>, grid, block, 0, st, l, vocab, stride, (int64_t*)token_out, (int64_t*)pos_inout, (int64_t*)history, history_len));
  else if (token_is_int64) B200_CUDA(launch(argmax_kernel<int64_t, int32_t>, grid, block, 0, st, l, vocab, stride, (int64_t*)token_out, (int32_t*)pos_inout, (int64_t*)history, history_len));
  else if (pos_is_int64) B200_CUDA(launch(argmax_kernel<int32_t, int64_t>, grid, block, 0, st, l, vocab, stride, (int32_t*)token_out, (int64_t*)pos_inout, (int32_t*)history, history_len));
  else B200_CUDA(launch(argmax_kernel<int32_t, int32_t>, grid, block, 0, st, l, vocab, stride, (int32_t*)token_out, (int32_t*)pos_inout, (int32_t*)history, history_len));
  return B200_OK;
}

}  // extern "C"
