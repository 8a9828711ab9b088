// Kernel-side "tiled" weight layout for the streaming dequant-GEMV, and its inverse.
//
// The reference layout (bitsandbytes Params4bit: row-major packed nibbles + a separate absmax vector; this
// project's int4: row-major qweight + {scale, zero} words) makes a warp that feeds mma.sync touch 8 different
// rows per load instruction.  The tiled layout stores, for every 16-row tile R and every 64-wide k-block kb,
// ONE contiguous record in exactly the order the 32 lanes of a warp consume it, so the whole weight stream of a
// warp is sequential and every load instruction is a fully coalesced 512 B (weights) or 64 B (scales) access:
//
//   NF4 / FP4 record (576 B), index R * nkb + kb:
//     [lane 0..31][16 B]   lane = 4*g + t:  bytes 0..7  = packed bytes [32*kb + 8*t, +8) of row 16*R + g
//                                            bytes 8..15 = the same bytes of row 16*R + g + 8
//     [g 0..7][8 B]        float absmax(row 16R+g, kb), float absmax(row 16R+g+8, kb)
//   INT4 record (1088 B), index R * (nkb/2) + kb/2  (group 128 = two k-blocks share one {scale, zero}):
//     [2 k-blocks][lane][16 B] as above, then [g][8 B] = half2{scale,zero}(row g), half2{scale,zero}(row g+8);
//     inside every 32-bit word the 8 codes e0..e7 are re-ordered to e0 e2 e4 e6 | e1 e3 e5 e7 (even elements in the
//     low half-word, odd ones in the high half-word) so that ((w >> 4p) & 0x000F000F) | 0x43004300 IS the bf16 pair
//     (128 + e_2p, 128 + e_2p+1): the GEMV dequantises int4 with one shift and one LOP3 per two weights, no table.
//
// Rows beyond N (N not a multiple of 16) are zero with scale 0.  The transform is a pure permutation of the
// reference bytes (int4: of the reference nibbles): b200_untile_4bit reproduces the reference arrays bit for bit
// (tests/test_gpu_kernels.py).
#include "common.cuh"

namespace b200 {

// int4 word: codes e_i at bits 4i  <->  even codes at bits 4p (p = 0..3), odd codes at bits 16 + 4p
__device__ __forceinline__ uint32_t int4_split_even_odd(uint32_t w) {
  uint32_t r = 0;
#pragma unroll
  for (int p = 0; p < 4; ++p) r |= (((w >> (8 * p)) & 15u) << (4 * p)) | (((w >> (8 * p + 4)) & 15u) << (16 + 4 * p));
  return r;
}
__device__ __forceinline__ uint32_t int4_merge_even_odd(uint32_t w) {
  uint32_t r = 0;
#pragma unroll
  for (int p = 0; p < 4; ++p) r |= (((w >> (4 * p)) & 15u) << (8 * p)) | (((w >> (16 + 4 * p)) & 15u) << (8 * p + 4));
  return r;
}

__global__ void __launch_bounds__(256) tile_4bit_kernel(const uint8_t* __restrict__ packed, const uint32_t* __restrict__ scale,
                                                        int64_t N, int64_t K, int int4, uint8_t* __restrict__ out, int64_t nrec_total) {
  griddep_wait();
  const int nkb = (int)(K / 64);
  const int bpr = int4 ? 2 : 1;                  // k-blocks per record
  const int rec_bytes = bpr * 512 + 64;
  const int nrec_row = nkb / bpr;                // records per row tile
  const int sc_per_row = int4 ? nkb / 2 : nkb;   // scale words per matrix row
  const int lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const int64_t warps_total = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t rec = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); rec < nrec_total; rec += warps_total) {
    const int64_t R = rec / nrec_row;
    const int j = (int)(rec % nrec_row);
    uint8_t* dst = out + rec * rec_bytes;
    const int64_t r0 = R * 16 + g, r1 = r0 + 8;
    for (int b = 0; b < bpr; ++b) {
      const int kb = j * bpr + b;
      uint2 lo = make_uint2(0u, 0u), hi = make_uint2(0u, 0u);
      if (r0 < N) lo = *reinterpret_cast<const uint2*>(packed + r0 * (K >> 1) + kb * 32 + t * 8);
      if (r1 < N) hi = *reinterpret_cast<const uint2*>(packed + r1 * (K >> 1) + kb * 32 + t * 8);
      if (int4) {
        lo.x = int4_split_even_odd(lo.x); lo.y = int4_split_even_odd(lo.y);
        hi.x = int4_split_even_odd(hi.x); hi.y = int4_split_even_odd(hi.y);
      }
      *reinterpret_cast<uint4*>(dst + b * 512 + lane * 16) = make_uint4(lo.x, lo.y, hi.x, hi.y);
    }
    if (t == 0) {
      uint32_t s0 = r0 < N ? scale[r0 * sc_per_row + j] : 0u;
      uint32_t s1 = r1 < N ? scale[r1 * sc_per_row + j] : 0u;
      *reinterpret_cast<uint2*>(dst + bpr * 512 + g * 8) = make_uint2(s0, s1);
    }
  }
}

__global__ void __launch_bounds__(256) untile_4bit_kernel(const uint8_t* __restrict__ tiled, int64_t N, int64_t K, int int4,
                                                          uint8_t* __restrict__ packed, uint32_t* __restrict__ scale, int64_t nrec_total) {
  griddep_wait();
  const int nkb = (int)(K / 64);
  const int bpr = int4 ? 2 : 1;
  const int rec_bytes = bpr * 512 + 64;
  const int nrec_row = nkb / bpr;
  const int sc_per_row = int4 ? nkb / 2 : nkb;
  const int lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const int64_t warps_total = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t rec = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); rec < nrec_total; rec += warps_total) {
    const int64_t R = rec / nrec_row;
    const int j = (int)(rec % nrec_row);
    const uint8_t* src = tiled + rec * rec_bytes;
    const int64_t r0 = R * 16 + g, r1 = r0 + 8;
    for (int b = 0; b < bpr; ++b) {
      const int kb = j * bpr + b;
      uint4 v = *reinterpret_cast<const uint4*>(src + b * 512 + lane * 16);
      if (int4) {
        v.x = int4_merge_even_odd(v.x); v.y = int4_merge_even_odd(v.y);
        v.z = int4_merge_even_odd(v.z); v.w = int4_merge_even_odd(v.w);
      }
      if (r0 < N) *reinterpret_cast<uint2*>(packed + r0 * (K >> 1) + kb * 32 + t * 8) = make_uint2(v.x, v.y);
      if (r1 < N) *reinterpret_cast<uint2*>(packed + r1 * (K >> 1) + kb * 32 + t * 8) = make_uint2(v.z, v.w);
    }
    if (t == 0) {
      const uint2 s = *reinterpret_cast<const uint2*>(src + bpr * 512 + g * 8);
      if (r0 < N) scale[r0 * sc_per_row + j] = s.x;
      if (r1 < N) scale[r1 * sc_per_row + j] = s.y;
    }
  }
}

static int tile_check(const void* a, const void* b, const void* c, int64_t N, int64_t K, int qtype, int blocksize, const char* who) {
  B200_CHECK_ARG(a && b && c, "%s: null pointer", who);
  B200_CHECK_ARG(N > 0 && K > 0, "%s: N and K must be positive", who);
  B200_CHECK_ARG(qtype == B200_Q_NF4 || qtype == B200_Q_FP4 || qtype == B200_Q_INT4, "%s: bad qtype %d", who, qtype);
  if (qtype == B200_Q_INT4)
    B200_CHECK_ARG(blocksize == 128 && K % 128 == 0, "%s: int4 tiling needs group 128 and K %% 128 == 0 (K=%lld group=%d)", who, (long long)K, blocksize);
  else
    B200_CHECK_ARG(blocksize == 64 && K % 64 == 0, "%s: 4-bit tiling needs blocksize 64 and K %% 64 == 0 (K=%lld blocksize=%d)", who, (long long)K, blocksize);
  B200_CHECK_ARG(((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(c)) & 15) == 0, "%s: buffers must be 16-byte aligned", who);
  return B200_OK;
}

}  // namespace b200

using namespace b200;

extern "C" {

size_t b200_tiled_bytes(int64_t N, int64_t K, int qtype) {
  if (N <= 0 || K <= 0 || K % 64 != 0) return 0;
  const int64_t tiles = (N + 15) / 16, nkb = K / 64;
  if (qtype == B200_Q_INT4) return (K % 128) ? 0 : (size_t)(tiles * (nkb / 2) * 1088);
  return (size_t)(tiles * nkb * 576);
}

int b200_tile_4bit(const void* packed, const void* scale, int64_t N, int64_t K, int qtype, int blocksize, void* tiled,
                   b200_stream_t stream) {
  int rc = tile_check(packed, scale, tiled, N, K, qtype, blocksize, "b200_tile_4bit");
  if (rc != B200_OK) return rc;
  const int int4 = qtype == B200_Q_INT4;
  const int64_t nrec = ((N + 15) / 16) * (K / 64) / (int4 ? 2 : 1);
  int64_t blocks = (nrec + 7) / 8, cap = (int64_t)sm_count() * 16;
  B200_CUDA(launch(tile_4bit_kernel, dim3((unsigned)(blocks < cap ? blocks : cap)), dim3(256), 0, (cudaStream_t)stream,
                   (const uint8_t*)packed, (const uint32_t*)scale, N, K, int4, (uint8_t*)tiled, nrec));
  return B200_OK;
}

int b200_untile_4bit(const void* tiled, int64_t N, int64_t K, int qtype, int blocksize, void* packed, void* scale,
                     b200_stream_t stream) {
  int rc = tile_check(packed, scale, tiled, N, K, qtype, blocksize, "b200_untile_4bit");
  if (rc != B200_OK) return rc;
  const int int4 = qtype == B200_Q_INT4;
  const int64_t nrec = ((N + 15) / 16) * (K / 64) / (int4 ? 2 : 1);
  int64_t blocks = (nrec + 7) / 8, cap = (int64_t)sm_count() * 16;
  B200_CUDA(launch(untile_4bit_kernel, dim3((unsigned)(blocks < cap ? blocks : cap)), dim3(256), 0, (cudaStream_t)stream,
                   (const uint8_t*)tiled, N, K, int4, (uint8_t*)packed, (uint32_t*)scale, nrec));
  return B200_OK;
}

}  // extern "C"
