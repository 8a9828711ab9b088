// bnb.int8 -- LLM.int8() inference Linear (bitsandbytes 0.41 nn.Linear8bitLt(has_fp16_weights=False, threshold=6.0),
// selected by `--quantize bnb.int8` at generate/base.py:105,129-134; semantics restated in oracle/quant_ref.py:310-345).
//
//   weight:      CB  = rint(W * 127 / absmax_row)  int8 (N, K),   SCB = absmax_row fp32 (N)
//   activation:  columns k with any |x[m,k]| >= threshold are OUTLIERS: they leave the int8 product and go through an
//                fp16 side product against the dequantised weight columns; the rest is quantised per row,
//                CA = rint(x * 127 / absmax_row), SCA = absmax_row
//   y = f16( int32(CA . CB^T) * (SCA * SCB / 127^2) )  [+ bias -> f16]  [+ f16(x_out . f16(CB_out * SCB / 127)^T) -> f16]
//
// The int8 x int8 -> int32 products run on the tensor cores (mma.sync m16n8k32 s8: exact, order-free), weights as the
// A operand loaded straight from HBM in fragment order (16 contiguous bytes of one row per lane; the K permutation
// this implies is applied to both operands).  Two mappings over the same fragments:
//   rows <= 16 (decode): one CTA per 16 weight rows, its 8 warps split K, int32 partials meet in shared memory;
//   rows  > 16 (prefill): a CTA holds 128 weight rows x 64 activation rows, each warp 16 x 64 over all of K.
#include "common.cuh"

namespace b200 {

constexpr int kI8Threads = 256;
constexpr int kI8Warps = 8;

__device__ __forceinline__ void imma16832(int (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// ---- weight quantisation: one CTA per output row ---------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) quantize_int8_rows_kernel(const T* __restrict__ w, int64_t K, int8_t* __restrict__ cb,
                                                                 float* __restrict__ scb) {
  __shared__ float s_max[8];
  const int64_t n = blockIdx.x;
  const T* row = w + n * K;
  float m = 0.f;
  for (int64_t k = threadIdx.x; k < K; k += 256) m = fmaxf(m, fabsf(load_as_float<T>(row, k)));
  m = warp_max(m);
  if ((threadIdx.x & 31) == 0) s_max[threadIdx.x >> 5] = m;
  __syncthreads();
  m = s_max[0];
#pragma unroll
  for (int i = 1; i < 8; ++i) m = fmaxf(m, s_max[i]);
  if (threadIdx.x == 0) scb[n] = m;
  const float inv = __fdiv_rn(127.0f, m);  // inf for an all-zero row: 0 * inf = NaN -> 0 below (np.nan_to_num)
  for (int64_t k = threadIdx.x; k < K; k += 256) {
    const float v = __fmul_rn(load_as_float<T>(row, k), inv);
    cb[n * K + k] = (v != v) ? (int8_t)0 : (int8_t)__float2int_rn(v);
  }
}

template <typename T>
__global__ void __launch_bounds__(256) dequantize_int8_rows_kernel(const int8_t* __restrict__ cb, const float* __restrict__ scb, int64_t N,
                                                                   int64_t K, T* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= N * K) return;
  store_from_float<T>(out, i, __fmul_rn((float)cb[i], __fdiv_rn(scb[i / K], 127.0f)));
}

// ---- activation side ------------------------------------------------------------------------------------------
// flags[k] = 1 when any row has |x[m,k]| >= threshold
__global__ void __launch_bounds__(256) int8_outlier_flags_kernel(const __half* __restrict__ x, int64_t M, int64_t K, float threshold,
                                                                 int* __restrict__ flags) {
  griddep_wait();
  const int64_t k = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (k >= K) return;
  int f = 0;
  if (threshold > 0.f) {
    for (int64_t m = 0; m < M; ++m) f |= fabsf(__half2float(x[m * K + k])) >= threshold;
  }
  flags[k] = f;
}

// ordered compaction of the flagged columns (ascending k: the fp32 side product is summed in this order)
__global__ void __launch_bounds__(1024) int8_compact_kernel(const int* __restrict__ flags, int64_t K, int* __restrict__ idx,
                                                            int* __restrict__ count) {
  __shared__ int s_cnt[1024];
  griddep_wait();
  const int t = threadIdx.x;
  const int64_t per = (K + 1023) / 1024, beg = t * per, end = beg + per < K ? beg + per : K;
  int c = 0;
  for (int64_t k = beg; k < end; ++k) c += flags[k];
  s_cnt[t] = c;
  __syncthreads();
  for (int o = 1; o < 1024; o <<= 1) {  // inclusive scan
    const int v = t >= o ? s_cnt[t - o] : 0;
    __syncthreads();
    s_cnt[t] += v;
    __syncthreads();
  }
  int pos = s_cnt[t] - c;
  for (int64_t k = beg; k < end; ++k) {
    if (flags[k]) idx[pos++] = (int)k;
  }
  if (t == 1023) *count = s_cnt[1023];
}

// one CTA per activation row: SCA = absmax over the non-outlier columns, CA = rint(x * 127 / SCA) (0 in outlier columns)
__global__ void __launch_bounds__(256) int8_rowquant_kernel(const __half* __restrict__ x, int64_t K, const int* __restrict__ flags,
                                                            int8_t* __restrict__ ca, float* __restrict__ sca) {
  __shared__ float s_max[8];
  griddep_wait();
  const int64_t m = blockIdx.x;
  const __half* row = x + m * K;
  float mx = 0.f;
  for (int64_t k = threadIdx.x; k < K; k += 256) mx = fmaxf(mx, flags[k] ? 0.f : fabsf(__half2float(row[k])));
  mx = warp_max(mx);
  if ((threadIdx.x & 31) == 0) s_max[threadIdx.x >> 5] = mx;
  __syncthreads();
  mx = s_max[0];
#pragma unroll
  for (int i = 1; i < 8; ++i) mx = fmaxf(mx, s_max[i]);
  if (threadIdx.x == 0) sca[m] = mx;
  const float inv = __fdiv_rn(127.0f, mx);
  for (int64_t k = threadIdx.x; k < K; k += 256) {
    const float v = __fmul_rn(flags[k] ? 0.f : __half2float(row[k]), inv);
    ca[m * K + k] = (v != v) ? (int8_t)0 : (int8_t)__float2int_rn(v);
  }
}

struct I8Params {
  const int8_t* cb;      // (N, K)
  const float* scb;      // (N)
  const __half* x;       // (M, K) original activations (outlier side product)
  const int8_t* ca;      // (M, K)
  const float* sca;      // (M)
  const int* idx;        // outlier columns, ascending
  const int* count;
  const __half* bias;    // (N) or null
  __half* y;             // (M, N)
  int64_t M, N, K;
};

__device__ __forceinline__ float round_f16(float v) { return __half2float(__float2half_rn(v)); }

// oracle/quant_ref.py:337-345: scale, round, bias, round, outlier side product, round
__device__ __forceinline__ void i8_store(const I8Params& p, int64_t m, int64_t n, int acc) {
  const float scale = __fdiv_rn(__fmul_rn(p.sca[m], p.scb[n]), 127.0f * 127.0f);
  float out = round_f16(__fmul_rn((float)acc, scale));
  if (p.bias) out = round_f16(out + __half2float(p.bias[n]));
  const int no = *p.count;
  if (no > 0) {
    const float wscale = __fdiv_rn(p.scb[n], 127.0f);
    float sub = 0.f;
    for (int j = 0; j < no; ++j) {
      const int k = p.idx[j];
      sub = fmaf(__half2float(p.x[m * p.K + k]), round_f16(__fmul_rn((float)p.cb[n * p.K + k], wscale)), sub);
    }
    out = round_f16(out + round_f16(sub));
  }
  p.y[m * p.N + n] = __float2half_rn(out);
}

// ---- decode: M <= 16.  grid = ceil(N / 16); 8 warps split K in 64-byte chunks ------------------------------------
template <int MB>
__global__ void __launch_bounds__(kI8Threads) int8_gemv_kernel(const I8Params p) {
  __shared__ int s_red[kI8Warps][16][8 * MB];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
  const int64_t n0 = (int64_t)blockIdx.x * 16;
  const int64_t ra = n0 + g < p.N ? n0 + g : p.N - 1, rb = n0 + g + 8 < p.N ? n0 + g + 8 : p.N - 1;
  const int8_t* wa = p.cb + ra * p.K + 16 * t;
  const int8_t* wb = p.cb + rb * p.K + 16 * t;
  const int nchunks = (int)(p.K >> 6);
  int acc[MB][4];
#pragma unroll
  for (int mb = 0; mb < MB; ++mb) acc[mb][0] = acc[mb][1] = acc[mb][2] = acc[mb][3] = 0;

  constexpr int U = 4;  // chunks in flight per warp
  uint4 a[U], b[U];
  // the weights do not depend on the previous kernel: request the first chunks before waiting for it
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const int c = warp + u * kI8Warps;
    a[u] = b[u] = make_uint4(0u, 0u, 0u, 0u);
    if (c < nchunks) { a[u] = ldg_stream_v4(wa + (int64_t)c * 64); b[u] = ldg_stream_v4(wb + (int64_t)c * 64); }
  }
  griddep_launch();
  griddep_wait();
  for (int c0 = warp; c0 < nchunks; c0 += U * kI8Warps) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int c = c0 + u * kI8Warps;
      if (c < nchunks) {
        const uint4 av = a[u], bv = b[u];
        const int cn = c + U * kI8Warps;
        if (cn < nchunks) { a[u] = ldg_stream_v4(wa + (int64_t)cn * 64); b[u] = ldg_stream_v4(wb + (int64_t)cn * 64); }
#pragma unroll
        for (int mb = 0; mb < MB; ++mb) {
          const int64_t m = mb * 8 + g;
          uint4 xv = make_uint4(0u, 0u, 0u, 0u);
          if (m < p.M) xv = *reinterpret_cast<const uint4*>(p.ca + m * p.K + (int64_t)c * 64 + 16 * t);
          imma16832(acc[mb], av.x, bv.x, av.y, bv.y, xv.x, xv.y);
          imma16832(acc[mb], av.z, bv.z, av.w, bv.w, xv.z, xv.w);
        }
      }
    }
  }
#pragma unroll
  for (int mb = 0; mb < MB; ++mb) {
    s_red[warp][g][8 * mb + 2 * t] = acc[mb][0];
    s_red[warp][g][8 * mb + 2 * t + 1] = acc[mb][1];
    s_red[warp][g + 8][8 * mb + 2 * t] = acc[mb][2];
    s_red[warp][g + 8][8 * mb + 2 * t + 1] = acc[mb][3];
  }
  __syncthreads();
  for (int o = tid; o < 16 * 8 * MB; o += kI8Threads) {
    const int r = o / (8 * MB), c = o % (8 * MB);
    if (c < p.M && n0 + r < p.N) {
      int s = 0;
#pragma unroll
      for (int w = 0; w < kI8Warps; ++w) s += s_red[w][r][c];
      i8_store(p, c, n0 + r, s);
    }
  }
}

// ---- prefill: M > 16.  grid = (ceil(N / 128), ceil(M / 64)); warp w: weight rows [n0 + 16 w, +16) x 64 activation rows
__global__ void __launch_bounds__(kI8Threads) int8_gemm_kernel(const I8Params p) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
  const int64_t n0 = (int64_t)blockIdx.x * 128 + warp * 16, m0 = (int64_t)blockIdx.y * 64;
  if (n0 >= p.N) return;
  const int64_t ra = n0 + g < p.N ? n0 + g : p.N - 1, rb = n0 + g + 8 < p.N ? n0 + g + 8 : p.N - 1;
  const int8_t* wa = p.cb + ra * p.K + 16 * t;
  const int8_t* wb = p.cb + rb * p.K + 16 * t;
  const int8_t* xr[8];
  bool xv_ok[8];
#pragma unroll
  for (int mb = 0; mb < 8; ++mb) {
    const int64_t m = m0 + mb * 8 + g;
    xv_ok[mb] = m < p.M;
    xr[mb] = p.ca + (xv_ok[mb] ? m : 0) * p.K + 16 * t;
  }
  const int nchunks = (int)(p.K >> 6);
  int acc[8][4];
#pragma unroll
  for (int mb = 0; mb < 8; ++mb) acc[mb][0] = acc[mb][1] = acc[mb][2] = acc[mb][3] = 0;
  griddep_launch();
  griddep_wait();
  uint4 an = *reinterpret_cast<const uint4*>(wa), bn = *reinterpret_cast<const uint4*>(wb);
  for (int c = 0; c < nchunks; ++c) {
    const uint4 av = an, bv = bn;
    if (c + 1 < nchunks) {
      an = *reinterpret_cast<const uint4*>(wa + (int64_t)(c + 1) * 64);
      bn = *reinterpret_cast<const uint4*>(wb + (int64_t)(c + 1) * 64);
    }
    uint4 xv[8];
#pragma unroll
    for (int mb = 0; mb < 8; ++mb) {
      xv[mb] = make_uint4(0u, 0u, 0u, 0u);
      if (xv_ok[mb]) xv[mb] = *reinterpret_cast<const uint4*>(xr[mb] + (int64_t)c * 64);
    }
#pragma unroll
    for (int mb = 0; mb < 8; ++mb) {
      imma16832(acc[mb], av.x, bv.x, av.y, bv.y, xv[mb].x, xv[mb].y);
      imma16832(acc[mb], av.z, bv.z, av.w, bv.w, xv[mb].z, xv[mb].w);
    }
  }
#pragma unroll
  for (int mb = 0; mb < 8; ++mb) {
    const int64_t m = m0 + mb * 8 + 2 * t;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int64_t mm = m + (i & 1), nn = n0 + g + (i >> 1) * 8;
      if (mm < p.M && nn < p.N) i8_store(p, mm, nn, acc[mb][i]);
    }
  }
}

}  // namespace b200

using namespace b200;

extern "C" {

int b200_quantize_int8(const void* w, int w_dtype, int64_t N, int64_t K, int8_t* cb, float* scb, b200_stream_t stream) {
  B200_CHECK_ARG(w && cb && scb, "b200_quantize_int8: null pointer");
  B200_CHECK_ARG(N > 0 && K > 0 && N <= 0x7fffffffLL, "b200_quantize_int8: bad shape (N=%lld K=%lld)", (long long)N, (long long)K);
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e;
  if (w_dtype == B200_BF16) e = launch(quantize_int8_rows_kernel<__nv_bfloat16>, dim3((unsigned)N), dim3(256), 0, st, (const __nv_bfloat16*)w, K, cb, scb);
  else if (w_dtype == B200_F16) e = launch(quantize_int8_rows_kernel<__half>, dim3((unsigned)N), dim3(256), 0, st, (const __half*)w, K, cb, scb);
  else if (w_dtype == B200_F32) e = launch(quantize_int8_rows_kernel<float>, dim3((unsigned)N), dim3(256), 0, st, (const float*)w, K, cb, scb);
  else { set_error("b200_quantize_int8: bad dtype code %d", w_dtype); return B200_ERR_INVALID; }
  if (e != cudaSuccess) { set_error("b200_quantize_int8: launch failed: %s", cudaGetErrorString(e)); return B200_ERR_CUDA; }
  return B200_OK;
}

int b200_dequantize_int8(const int8_t* cb, const float* scb, int64_t N, int64_t K, void* out, int out_dtype, b200_stream_t stream) {
  B200_CHECK_ARG(cb && scb && out, "b200_dequantize_int8: null pointer");
  B200_CHECK_ARG(N > 0 && K > 0, "b200_dequantize_int8: bad shape");
  cudaStream_t st = (cudaStream_t)stream;
  const dim3 grid((unsigned)((N * K + 255) / 256));
  cudaError_t e;
  if (out_dtype == B200_BF16) e = launch(dequantize_int8_rows_kernel<__nv_bfloat16>, grid, dim3(256), 0, st, cb, scb, N, K, (__nv_bfloat16*)out);
  else if (out_dtype == B200_F16) e = launch(dequantize_int8_rows_kernel<__half>, grid, dim3(256), 0, st, cb, scb, N, K, (__half*)out);
  else if (out_dtype == B200_F32) e = launch(dequantize_int8_rows_kernel<float>, grid, dim3(256), 0, st, cb, scb, N, K, (float*)out);
  else { set_error("b200_dequantize_int8: bad dtype code %d", out_dtype); return B200_ERR_INVALID; }
  if (e != cudaSuccess) { set_error("b200_dequantize_int8: launch failed: %s", cudaGetErrorString(e)); return B200_ERR_CUDA; }
  return B200_OK;
}

size_t b200_linear_int8_workspace_bytes(int64_t M, int64_t K) {
  // flags int32[K] | idx int32[K] | count (16 B) | sca fp32[M] (padded to 16) | ca int8[M * K]
  return (size_t)K * 8 + 16 + (((size_t)M * 4 + 15) & ~(size_t)15) + (size_t)M * K;
}

int b200_linear_int8(const b200_linear_int8_args* a, b200_stream_t stream) {
  B200_CHECK_ARG(a != nullptr, "b200_linear_int8: null args");
  B200_CHECK_ARG(a->x && a->cb && a->scb && a->y && a->workspace, "b200_linear_int8: null x/cb/scb/y/workspace pointer");
  B200_CHECK_ARG(a->M > 0 && a->N > 0 && a->K > 0, "b200_linear_int8: M, N, K must be positive");
  B200_CHECK_ARG(a->workspace_bytes >= b200_linear_int8_workspace_bytes(a->M, a->K), "b200_linear_int8: workspace too small (%zu < %zu)",
                 a->workspace_bytes, b200_linear_int8_workspace_bytes(a->M, a->K));
  B200_CHECK_ARG(((reinterpret_cast<uintptr_t>(a->x) | reinterpret_cast<uintptr_t>(a->cb) | reinterpret_cast<uintptr_t>(a->workspace)) & 15) == 0,
                 "b200_linear_int8: x / cb / workspace must be 16-byte aligned");
  if (a->K % 64 != 0) {
    set_error("b200_linear_int8: K must be a multiple of 64 (got %lld)", (long long)a->K);
    return B200_ERR_UNSUPPORTED;
  }
  B200_CHECK_ARG((a->M + 63) / 64 <= 65535, "b200_linear_int8: too many activation rows for the grid");
  uint8_t* ws = reinterpret_cast<uint8_t*>(a->workspace);
  int* flags = reinterpret_cast<int*>(ws);
  int* idx = flags + a->K;
  int* count = idx + a->K;
  float* sca = reinterpret_cast<float*>(ws + (size_t)a->K * 8 + 16);
  int8_t* ca = reinterpret_cast<int8_t*>(ws + (size_t)a->K * 8 + 16 + (((size_t)a->M * 4 + 15) & ~(size_t)15));
  const __half* x = reinterpret_cast<const __half*>(a->x);
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e = launch(int8_outlier_flags_kernel, dim3((unsigned)((a->K + 255) / 256)), dim3(256), 0, st, x, a->M, a->K, a->threshold, flags);
  if (e == cudaSuccess) e = launch(int8_compact_kernel, dim3(1), dim3(1024), 0, st, (const int*)flags, a->K, idx, count);
  if (e == cudaSuccess) e = launch(int8_rowquant_kernel, dim3((unsigned)a->M), dim3(256), 0, st, x, a->K, (const int*)flags, ca, sca);
  I8Params p{};
  p.cb = reinterpret_cast<const int8_t*>(a->cb);
  p.scb = reinterpret_cast<const float*>(a->scb);
  p.x = x; p.ca = ca; p.sca = sca; p.idx = idx; p.count = count;
  p.bias = reinterpret_cast<const __half*>(a->bias);
  p.y = reinterpret_cast<__half*>(a->y);
  p.M = a->M; p.N = a->N; p.K = a->K;
  if (e == cudaSuccess) {
    const dim3 gv((unsigned)((a->N + 15) / 16));
    if (a->M <= 8) e = launch(int8_gemv_kernel<1>, gv, dim3(kI8Threads), 0, st, p);
    else if (a->M <= 16) e = launch(int8_gemv_kernel<2>, gv, dim3(kI8Threads), 0, st, p);
    else e = launch(int8_gemm_kernel, dim3((unsigned)((a->N + 127) / 128), (unsigned)((a->M + 63) / 64)), dim3(kI8Threads), 0, st, p);
  }
  if (e != cudaSuccess) {
    set_error("b200_linear_int8: launch failed: %s", cudaGetErrorString(e));
    return B200_ERR_CUDA;
  }
  return B200_OK;
}

}  // extern "C"
