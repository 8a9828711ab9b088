// Shared device/host helpers for the b200gpt kernels (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "../../include/b200gpt.h"

namespace b200 {

// ---- host: errors, launch bookkeeping -------------------------------------------------------------------
void set_error(const char* fmt, ...);
extern std::atomic<int64_t> g_launches;
extern std::atomic<int> g_pdl;

#define B200_CHECK_ARG(cond, ...)        \
  do {                                   \
    if (!(cond)) {                       \
      b200::set_error(__VA_ARGS__);      \
      return B200_ERR_INVALID;           \
    }                                    \
  } while (0)

#define B200_CUDA(expr)                                                                       \
  do {                                                                                        \
    cudaError_t _e = (expr);                                                                  \
    if (_e != cudaSuccess) {                                                                  \
      b200::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return B200_ERR_CUDA;                                                                   \
    }                                                                                         \
  } while (0)

// Launch with the programmatic-dependent-launch attribute when enabled: the kernel may start while its
// predecessor in the stream drains; it must execute griddep_wait() before touching the predecessor's outputs
// and before its first global write.
template <typename... KArgs, typename... Args>
inline cudaError_t launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = g_pdl.load(std::memory_order_relaxed) ? 1 : 0;
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

int sm_count();

// Kernel attributes (cudaFuncSetAttribute) are per DEVICE: remember on which devices an instantiation has been configured, so that
// one process driving several GPUs (generate/sequentially.py) configures each of them.
struct PerDeviceOnce {
  std::atomic<unsigned long long> mask{0};
  bool need() const { int d = 0; cudaGetDevice(&d); return ((mask.load(std::memory_order_acquire) >> (d & 63)) & 1ull) == 0; }
  void done() { int d = 0; cudaGetDevice(&d); mask.fetch_or(1ull << (d & 63), std::memory_order_release); }
};

// ---- device helpers ---------------------------------------------------------------------------------------
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void griddep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

__device__ __forceinline__ float bf16_bits_to_float(uint32_t lo16) { return __uint_as_float(lo16 << 16); }
__device__ __forceinline__ float bf16lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf16hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 p = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&p);
}
__device__ __forceinline__ float round_bf16(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }
__device__ __forceinline__ uint16_t bf16_bits(float v) {
  __nv_bfloat16 h = __float2bfloat16_rn(v);
  return *reinterpret_cast<uint16_t*>(&h);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// int4 zero-point term: xsum[b * nkb + kb] = sum of the 64 bf16 activations of k-block kb of row b (rows in shared memory,
// `xstride` elements apart).  Eight lanes per k-block, each reading four words whose order is rotated by the k-block's
// slot so that the four k-blocks of a warp hit distinct banks (the one-thread-per-k-block loop this replaces had every lane
// of a warp on the same bank: 32-way conflicts, ~1 us per launch), then a fixed shuffle tree: deterministic.
// All NT threads must call it (NT a multiple of 32); the caller barriers afterwards.
template <int NT>
__device__ __forceinline__ void xsum_blocks(const __nv_bfloat16* x_s, int xstride, int B, int nkb, float* xsum_s, int tid) {
  const int total = B * nkb, sub = tid & 7, slot = tid >> 3;
  for (int base = 0; base < total; base += NT / 8) {
    const int i = base + slot;
    float s = 0.f;
    if (i < total) {
      const int b = i / nkb, kb = i % nkb;
      const uint32_t* px = reinterpret_cast<const uint32_t*>(x_s + b * xstride + kb * 64);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const uint32_t w = px[sub + 8 * ((j + slot) & 3)];
        s += bf16lo(w) + bf16hi(w);
      }
    }
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    s += __shfl_xor_sync(0xffffffffu, s, 4);
    if (sub == 0 && i < total) xsum_s[i] = s;
  }
}

// streaming 128-bit / 64-bit loads that bypass L1 allocation (weights are read exactly once)
__device__ __forceinline__ uint4 ldg_stream_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ uint2 ldg_stream_v2(const void* p) {
  uint2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ uint32_t ldg_stream_u32(const void* p) {
  uint32_t r;
  asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
  return r;
}

// 4-bit codebooks (bitsandbytes 0.41 kernels.cu dDequantizeNF4 / dDequantizeFP4Tree)
__device__ __constant__ const float kNF4[16] = {
    -1.0f, -0.6961928009986877f, -0.5250730514526367f, -0.39491748809814453f, -0.28444138169288635f,
    -0.18477343022823334f, -0.09105003625154495f, 0.0f, 0.07958029955625534f, 0.16093020141124725f,
    0.24611230194568634f, 0.33791524171829224f, 0.44070982933044434f, 0.5626170039176941f,
    0.7229568362236023f, 1.0f};
__device__ __constant__ const float kFP4[16] = {
    0.0f, 5.208333333e-03f, 0.66666667f, 1.0f, 0.33333333f, 0.5f, 0.16666667f, 0.25f,
    -0.0f, -5.208333333e-03f, -0.66666667f, -1.0f, -0.33333333f, -0.5f, -0.16666667f, -0.25f};

__device__ __forceinline__ float code_value(int qtype, int c) { return qtype == B200_Q_NF4 ? kNF4[c] : kFP4[c]; }

template <typename T>
__device__ __forceinline__ float load_as_float(const T* p, int64_t i);
template <>
__device__ __forceinline__ float load_as_float<__nv_bfloat16>(const __nv_bfloat16* p, int64_t i) { return __bfloat162float(p[i]); }
template <>
__device__ __forceinline__ float load_as_float<__half>(const __half* p, int64_t i) { return __half2float(p[i]); }
template <>
__device__ __forceinline__ float load_as_float<float>(const float* p, int64_t i) { return p[i]; }

template <typename T>
__device__ __forceinline__ void store_from_float(T* p, int64_t i, float v);
template <>
__device__ __forceinline__ void store_from_float<__nv_bfloat16>(__nv_bfloat16* p, int64_t i, float v) { p[i] = __float2bfloat16_rn(v); }
template <>
__device__ __forceinline__ void store_from_float<__half>(__half* p, int64_t i, float v) { p[i] = __float2half_rn(v); }
template <>
__device__ __forceinline__ void store_from_float<float>(float* p, int64_t i, float v) { p[i] = v; }

}  // namespace b200
