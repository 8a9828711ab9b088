// Fused dequant-GEMV for batch 1..16 decode (the dominant kernel of the hot path).
//
// Replaces bnb.nn.Linear4bit.forward at lit_gpt/model.py:619,656,713,714,716,519 (reference: the
// kgemm_4bit_inference_naive kernel for a single token, dequantize_4bit + cuBLAS otherwise).
//
// HBM-bound: every packed weight byte is read exactly once with coalesced streaming loads
// (ld.global.nc.L1::no_allocate), straight into registers, double-buffered one row-tile ahead.
// The issue-slot budget at 6.5 TB/s is ~3 warp-instructions per weight, so the dequant is:
//   * one PRMT per packed BYTE builds a shared-memory address  (byte << 8 | lane << 2),
//   * one conflict-free LDS (each lane owns a bank) from a 256-entry byte->bf16x2 codebook table returns TWO
//     dequantised weights already in mma.sync A-fragment order,
//   * one mma.sync.m16n8k16 (bf16 x bf16 -> fp32) per 256 weights does the multiply-accumulate against the
//     activations, which live in registers as B fragments for the whole kernel (batch = the n dimension),
//   * the per-64 absmax (or int4 group scale/zero) is applied in fp32 to the 4 accumulators after every 4 MMAs.
// K is split over the 8 warps of a CTA and over gridDim.y CTAs (deterministic "last arriver reduces" through
// a caller-owned workspace), so a warp's activation slice stays in <= 32 registers and small matrices still
// fill 148 SMs.  Row tiles are walked persistently; all global writes happen after griddepcontrol.wait, so the
// kernel's table init and first weight prefetch overlap the previous kernel's tail (PDL).
// Prologue (optional): residual add + RMSNorm of the activation (lit_gpt/rmsnorm.py:19-25, model.py:591).
// Epilogue: bias / residual add / SwiGLU (lit_gpt/model.py:712-716), with the reference's bf16 rounding points.
#include "common.cuh"

namespace b200 {

struct GemvParams {
  const uint8_t* wA;   // packed matrix A (rows g)
  const uint8_t* wB;   // packed matrix for rows g+8 (A + 8 rows, or the second matrix for SwiGLU)
  const uint32_t* sA;  // per-block scale words (float absmax for NF4/FP4, half2{scale,zero} for INT4)
  const uint32_t* sB;
  int64_t N, K;
  int nkb;             // K / 64
  int sc_per_row;      // scale words per row (K/64 for 4-bit block-64, K/128 for int4)
  int sc_shift;        // kb >> sc_shift = scale word index within the row (0 or 1)
  int rows_per_tile;   // 16, or 8 for SwiGLU
  int ntiles;
  int qtype;
  const __nv_bfloat16* x;
  const __nv_bfloat16* delta;
  __nv_bfloat16* x_out;
  const __nv_bfloat16* norm_w;
  float eps;
  int batch;
  int epilogue;
  const __nv_bfloat16* bias;
  const __nv_bfloat16* bias2;
  const __nv_bfloat16* residual;
  __nv_bfloat16* y;
  float* ws_part;      // [ntiles][S][16*COLS]
  unsigned int* ws_cnt;  // [ntiles]
};

constexpr int kGemvThreads = 256;
constexpr int kGemvWarps = 8;
constexpr int kLutBytes = 256 * 256;  // entry e, lane l at byte e*256 + l*4 (upper 128 B of each row unused)

__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                               uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

__device__ __forceinline__ float silu_bf16(float a) {  // F.silu on a bf16 tensor: fp32 math, one rounding
  return round_bf16(a / (1.0f + expf(-a)));
}

// value the epilogue writes for one (row, col); h is the fp32 accumulator
__device__ __forceinline__ void epilogue_store(const GemvParams& p, int64_t n, int c, float vA, float vB) {
  if (p.epilogue == B200_EPI_SWIGLU) {
    float a = round_bf16(vA + (p.bias ? __bfloat162float(p.bias[n]) : 0.f));
    float b = round_bf16(vB + (p.bias2 ? __bfloat162float(p.bias2[n]) : 0.f));
    p.y[(int64_t)c * p.N + n] = __float2bfloat16_rn(silu_bf16(a) * b);
    return;
  }
  float h = vA + (p.bias ? __bfloat162float(p.bias[n]) : 0.f);
  if (p.epilogue == B200_EPI_RESIDUAL) h = round_bf16(h) + __bfloat162float(p.residual[(int64_t)c * p.N + n]);
  p.y[(int64_t)c * p.N + n] = __float2bfloat16_rn(h);
}

// ------------------------------------------------------------------------------------------------------------
// Generic kernel: any N, K, blocksize, batch <= 16; one warp per output row; scalar nibble decode.
// Used for shapes the fast path does not cover (K % 64 != 0, toy models).  Activation prologue is done by a
// separate rmsnorm launch into the workspace, so x here is already final.
// ------------------------------------------------------------------------------------------------------------
struct GenericParams {
  const uint8_t* w[2];
  const uint32_t* s[2];
  int64_t N, K;
  int blocksize, qtype, batch, epilogue;
  const __nv_bfloat16* x;
  const __nv_bfloat16* bias;
  const __nv_bfloat16* bias2;
  const __nv_bfloat16* residual;
  __nv_bfloat16* y;
};

__device__ __forceinline__ float generic_weight(const GenericParams& p, int m, int64_t n, int64_t k) {
  if (p.qtype == B200_Q_INT4) {
    const int64_t W8 = (p.K + 7) / 8, G = (p.K + p.blocksize - 1) / p.blocksize;
    const uint32_t* qw = reinterpret_cast<const uint32_t*>(p.w[m]);
    float q = (float)((qw[n * W8 + k / 8] >> (4 * (k % 8))) & 0xF);
    uint32_t sz = p.s[m][n * G + k / p.blocksize];
    float s = __half2float(__ushort_as_half((unsigned short)(sz & 0xffff)));
    float z = __half2float(__ushort_as_half((unsigned short)(sz >> 16)));
    return round_bf16(s * (q - z));
  }
  const int64_t e = n * p.K + k;
  unsigned char b = p.w[m][e >> 1];
  int c = (e & 1) ? (b & 15) : (b >> 4);
  return round_bf16(code_value(p.qtype, c) * __uint_as_float(p.s[m][e / p.blocksize]));
}

__global__ void __launch_bounds__(256) qgemv_generic_kernel(GenericParams p) {
  __shared__ float s_code[16];
  if (threadIdx.x < 16 && p.qtype != B200_Q_INT4) s_code[threadIdx.x] = code_value(p.qtype, threadIdx.x);
  griddep_wait();
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int64_t warps_total = (int64_t)gridDim.x * (blockDim.x >> 5);
  const int nmat = p.epilogue == B200_EPI_SWIGLU ? 2 : 1;
  // 4-bit code books with K a multiple of 8 (e.g. a row-parallel tensor-parallel shard whose K is not a multiple of the
  // 64-element block, so that blocks straddle rows and the tiled layout does not apply): eight weights per lane and
  // step -- one 32-bit load of packed codes, one absmax (eight aligned elements never cross a block), one 128-bit load of
  // activations -- instead of one byte load per weight.  Same value per weight as below: bf16(code * absmax).
  const bool vec8 = p.qtype != B200_Q_INT4 && nmat == 1 && (p.K & 7) == 0 && (p.blocksize & 7) == 0 &&
                    ((reinterpret_cast<uintptr_t>(p.x) & 15) == 0) && ((reinterpret_cast<uintptr_t>(p.w[0]) & 3) == 0);
  if (vec8) {
    for (int64_t n = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); n < p.N; n += warps_total) {
      float acc[16];
#pragma unroll
      for (int bb = 0; bb < 16; ++bb) acc[bb] = 0.f;
      for (int64_t k = (int64_t)lane * 8; k < p.K; k += 256) {
        const int64_t e = n * p.K + k;
        const uint32_t wq = __ldg(reinterpret_cast<const uint32_t*>(p.w[0] + (e >> 1)));
        const float am = __uint_as_float(__ldg(p.s[0] + e / p.blocksize));
        float w[8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {  // byte j holds elements 2j (high nibble) and 2j + 1 (low nibble)
          const uint32_t by = (wq >> (8 * j)) & 0xFFu;
          w[2 * j] = round_bf16(s_code[by >> 4] * am);
          w[2 * j + 1] = round_bf16(s_code[by & 15] * am);
        }
#pragma unroll
        for (int bb = 0; bb < 16; ++bb) {
          if (bb >= p.batch) break;
          const uint4 xv = *reinterpret_cast<const uint4*>(p.x + (int64_t)bb * p.K + k);
          const uint32_t xw[4] = {xv.x, xv.y, xv.z, xv.w};
          float a = acc[bb];
#pragma unroll
          for (int j = 0; j < 4; ++j) { a = fmaf(w[2 * j], bf16lo(xw[j]), a); a = fmaf(w[2 * j + 1], bf16hi(xw[j]), a); }
          acc[bb] = a;
        }
      }
#pragma unroll
      for (int bb = 0; bb < 16; ++bb) {
        if (bb >= p.batch) break;
        const float v = warp_sum(acc[bb]);
        if (lane == 0) {
          GemvParams q{};
          q.epilogue = p.epilogue; q.bias = p.bias; q.bias2 = p.bias2; q.residual = p.residual; q.y = p.y; q.N = p.N;
          epilogue_store(q, n, bb, v, 0.f);
        }
      }
    }
    return;
  }
  for (int64_t n = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); n < p.N; n += warps_total) {
    for (int b0 = 0; b0 < p.batch; b0 += 4) {
      float acc[2][4] = {};
      for (int64_t k = lane; k < p.K; k += 32) {
        float w0 = generic_weight(p, 0, n, k);
        float w1 = nmat == 2 ? generic_weight(p, 1, n, k) : 0.f;
#pragma unroll
        for (int bb = 0; bb < 4; ++bb) {
          if (b0 + bb < p.batch) {
            float xv = __bfloat162float(p.x[(int64_t)(b0 + bb) * p.K + k]);
            acc[0][bb] = fmaf(w0, xv, acc[0][bb]);
            acc[1][bb] = fmaf(w1, xv, acc[1][bb]);
          }
        }
      }
#pragma unroll
      for (int bb = 0; bb < 4; ++bb) {
        float vA = warp_sum(acc[0][bb]), vB = warp_sum(acc[1][bb]);
        if (lane == 0 && b0 + bb < p.batch) {
          GemvParams q{};
          q.epilogue = p.epilogue; q.bias = p.bias; q.bias2 = p.bias2; q.residual = p.residual; q.y = p.y; q.N = p.N;
          epilogue_store(q, n, b0 + bb, vA, vB);
        }
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------------------
// Fast kernel.  Requirements (checked by the host): K % 64 == 0, 4-bit block-64 or int4 group-128/256,
// batch <= 8*NB.  256 threads = 8 warps; warp w of split s owns k-blocks [kb0, kb0+cnt), cnt <= KS.
// Lane (g = lane/4, t = lane%4) owns, per k-block, 8 packed bytes (16 weights, k = 64*kb + 16*t + 0..15) of
// tile rows g and g+8, and the matching 16 activations of batch column g as mma B fragments.
// ------------------------------------------------------------------------------------------------------------
template <int KS>
struct WBuf {
  uint2 a[KS];      // 8 packed bytes of row g
  uint2 b[KS];      // 8 packed bytes of row g+8 (or row g of the second matrix)
  uint32_t sa[KS];  // scale word of row g for this k-block
  uint32_t sb[KS];
};

template <int KS>
__device__ __forceinline__ void load_tile(WBuf<KS>& buf, const GemvParams& p, int tile, int g, int t, int kb0, int cnt) {
  if (tile >= p.ntiles) return;
  const int64_t n0 = (int64_t)tile * p.rows_per_tile;
  int64_t rA = n0 + g, rB = (p.rows_per_tile == 16) ? n0 + 8 + g : n0 + g;
  rA = rA < p.N ? rA : p.N - 1;  // tail rows: load something valid, the result is never stored
  rB = rB < p.N ? rB : p.N - 1;
  const uint8_t* pa = p.wA + rA * (p.K >> 1) + (int64_t)kb0 * 32 + t * 8;
  const uint8_t* pb = p.wB + rB * (p.K >> 1) + (int64_t)kb0 * 32 + t * 8;
  const uint32_t* sa = p.sA + rA * p.sc_per_row;
  const uint32_t* sb = p.sB + rB * p.sc_per_row;
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
    if (ks < cnt) {
      buf.a[ks] = ldg_stream_v2(pa + ks * 32);
      buf.b[ks] = ldg_stream_v2(pb + ks * 32);
      buf.sa[ks] = ldg_stream_u32(sa + ((kb0 + ks) >> p.sc_shift));
      buf.sb[ks] = ldg_stream_u32(sb + ((kb0 + ks) >> p.sc_shift));
    }
  }
}

template <int KS, int NB, bool INT4>
__global__ void __launch_bounds__(kGemvThreads, 2) qgemv_fast_kernel(const GemvParams p) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  uint32_t* lut = reinterpret_cast<uint32_t*>(smem_raw);
  constexpr int COLS = 8 * NB;
  float* red = reinterpret_cast<float*>(smem_raw + kLutBytes);          // [2][8 warps][16][COLS]
  float* s_ssq = red + 2 * kGemvWarps * 16 * COLS;                      // [8 warps][16]
  float* s_rstd = s_ssq + kGemvWarps * 16;                              // [16]
  unsigned int* s_flag = reinterpret_cast<unsigned int*>(s_rstd + 16);  // [2]

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int S = gridDim.y, s = blockIdx.y;

  // ---- input-independent prologue (overlaps the previous kernel under PDL) -------------------------------
  for (int i = tid; i < 256 * 32; i += kGemvThreads) {
    const int e = i >> 5, l = i & 31;
    float lo, hi;
    if (INT4) { lo = (float)(e & 15); hi = (float)(e >> 4); }                      // k even = LOW nibble
    else { lo = code_value(p.qtype, e >> 4); hi = code_value(p.qtype, e & 15); }   // even element = HIGH nibble
    lut[e * 64 + l] = pack_bf16(lo, hi);
  }
  const int wg = s * kGemvWarps + warp, WT = S * kGemvWarps;
  const int kb0 = (int)(((int64_t)wg * p.nkb) / WT);
  const int cnt = (int)(((int64_t)(wg + 1) * p.nkb) / WT) - kb0;  // <= KS by host construction

  WBuf<KS> bufA, bufB;
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
    bufA.a[ks] = bufA.b[ks] = bufB.a[ks] = bufB.b[ks] = make_uint2(0u, 0u);
    bufA.sa[ks] = bufA.sb[ks] = bufB.sa[ks] = bufB.sb[ks] = 0u;
  }
  const int stride = gridDim.x;
  int tile = blockIdx.x;
  load_tile<KS>(bufA, p, tile, g, t, kb0, cnt);            // weights never depend on the previous kernel
  load_tile<KS>(bufB, p, tile + stride, g, t, kb0, cnt);
  griddep_launch();
  griddep_wait();

  // ---- activation prologue: (x + delta) -> optional RMSNorm -> B fragments in registers ------------------
  const int B = p.batch;
  const bool designated = (blockIdx.x == 0 && blockIdx.y == 0);
  if (p.norm_w != nullptr || (p.x_out != nullptr && designated)) {
    const int chunks = (int)(p.K >> 3);
    for (int b = 0; b < B; ++b) {
      float ss = 0.f;
      for (int c = tid; c < chunks; c += kGemvThreads) {
        uint4 v = *reinterpret_cast<const uint4*>(p.x + (int64_t)b * p.K + c * 8);
        uint32_t w4[4] = {v.x, v.y, v.z, v.w};
        if (p.delta) {
          uint4 d = *reinterpret_cast<const uint4*>(p.delta + (int64_t)b * p.K + c * 8);
          uint32_t d4[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
          for (int i = 0; i < 4; ++i) w4[i] = pack_bf16(bf16lo(w4[i]) + bf16lo(d4[i]), bf16hi(w4[i]) + bf16hi(d4[i]));
          if (p.x_out && designated)
            *reinterpret_cast<uint4*>(p.x_out + (int64_t)b * p.K + c * 8) = make_uint4(w4[0], w4[1], w4[2], w4[3]);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) { float a = bf16lo(w4[i]), c2 = bf16hi(w4[i]); ss = fmaf(a, a, ss); ss = fmaf(c2, c2, ss); }
      }
      ss = warp_sum(ss);
      if (lane == 0) s_ssq[warp * 16 + b] = ss;
    }
    __syncthreads();
    if (tid < B) {
      float tot = 0.f;
#pragma unroll
      for (int w = 0; w < kGemvWarps; ++w) tot += s_ssq[w * 16 + tid];
      s_rstd[tid] = rsqrtf(tot / (float)p.K + p.eps);
    }
  }
  __syncthreads();  // also publishes the LUT

  uint32_t xf[NB][KS][8];
  float xs[NB][KS][2];  // INT4 only: sum of the activations of this k-block for columns 2t, 2t+1
#pragma unroll
  for (int nb = 0; nb < NB; ++nb) {
    const int b = g + 8 * nb;
    const float rstd = (p.norm_w != nullptr && b < B) ? s_rstd[b] : 1.f;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      float lsum = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) xf[nb][ks][i] = 0u;
      if (b < B && ks < cnt) {
        const int64_t k0 = (int64_t)(kb0 + ks) * 64 + t * 16;
        const uint4* px = reinterpret_cast<const uint4*>(p.x + (int64_t)b * p.K + k0);
        uint4 v0 = px[0], v1 = px[1];
        uint32_t w8[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
        if (p.delta) {
          const uint4* pd = reinterpret_cast<const uint4*>(p.delta + (int64_t)b * p.K + k0);
          uint4 d0 = pd[0], d1 = pd[1];
          uint32_t d8[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
#pragma unroll
          for (int i = 0; i < 8; ++i) w8[i] = pack_bf16(bf16lo(w8[i]) + bf16lo(d8[i]), bf16hi(w8[i]) + bf16hi(d8[i]));
        }
        if (p.norm_w) {
          const uint4* pw = reinterpret_cast<const uint4*>(p.norm_w + k0);
          uint4 n0 = pw[0], n1 = pw[1];
          uint32_t n8[8] = {n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, n1.z, n1.w};
#pragma unroll
          for (int i = 0; i < 8; ++i)  // weight * (x * rstd), fp32, one cast (rmsnorm.py:24-25)
            w8[i] = pack_bf16(bf16lo(n8[i]) * (bf16lo(w8[i]) * rstd), bf16hi(n8[i]) * (bf16hi(w8[i]) * rstd));
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) { xf[nb][ks][i] = w8[i]; lsum += bf16lo(w8[i]) + bf16hi(w8[i]); }
      }
      if (INT4) {
        lsum += __shfl_xor_sync(0xffffffffu, lsum, 1);
        lsum += __shfl_xor_sync(0xffffffffu, lsum, 2);  // all 4 lanes of group g: sum over the 64 k of column g
        xs[nb][ks][0] = __shfl_sync(0xffffffffu, lsum, (2 * t) * 4);
        xs[nb][ks][1] = __shfl_sync(0xffffffffu, lsum, (2 * t + 1) * 4);
      } else {
        xs[nb][ks][0] = xs[nb][ks][1] = 0.f;
      }
    }
  }

  const uint32_t lane_off = (uint32_t)lane << 2;
  const uint8_t* lut_b = reinterpret_cast<const uint8_t*>(lut);
  auto LUT = [&](uint32_t word, uint32_t sel) -> uint32_t {
    // result = (byte_i << 8) | (lane << 2): PRMT picks lane_off.b0, word.b_i, 0, 0
    return *reinterpret_cast<const uint32_t*>(lut_b + __byte_perm(word, lane_off, sel));
  };

  auto process = [&](const WBuf<KS>& buf, int tl, int it) {
    float acc[NB][4];
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) acc[nb][0] = acc[nb][1] = acc[nb][2] = acc[nb][3] = 0.f;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      if (ks < cnt) {
        float part[NB][4];
#pragma unroll
        for (int nb = 0; nb < NB; ++nb) part[nb][0] = part[nb][1] = part[nb][2] = part[nb][3] = 0.f;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t wa = (j < 2) ? buf.a[ks].x : buf.a[ks].y;
          const uint32_t wb = (j < 2) ? buf.b[ks].x : buf.b[ks].y;
          const uint32_t sel0 = (j & 1) ? 0x5524u : 0x5504u, sel1 = (j & 1) ? 0x5534u : 0x5514u;
          const uint32_t a0 = LUT(wa, sel0), a2 = LUT(wa, sel1);
          const uint32_t a1 = LUT(wb, sel0), a3 = LUT(wb, sel1);
#pragma unroll
          for (int nb = 0; nb < NB; ++nb) mma_bf16_16816(part[nb], a0, a1, a2, a3, xf[nb][ks][2 * j], xf[nb][ks][2 * j + 1]);
        }
        if (INT4) {
          const __half2 ha = *reinterpret_cast<const __half2*>(&buf.sa[ks]);
          const __half2 hb = *reinterpret_cast<const __half2*>(&buf.sb[ks]);
          const float sa = __low2float(ha), za = __high2float(ha), sb = __low2float(hb), zb = __high2float(hb);
#pragma unroll
          for (int nb = 0; nb < NB; ++nb) {
            acc[nb][0] = fmaf(sa, part[nb][0] - za * xs[nb][ks][0], acc[nb][0]);
            acc[nb][1] = fmaf(sa, part[nb][1] - za * xs[nb][ks][1], acc[nb][1]);
            acc[nb][2] = fmaf(sb, part[nb][2] - zb * xs[nb][ks][0], acc[nb][2]);
            acc[nb][3] = fmaf(sb, part[nb][3] - zb * xs[nb][ks][1], acc[nb][3]);
          }
        } else {
          const float sa = __uint_as_float(buf.sa[ks]), sb = __uint_as_float(buf.sb[ks]);
#pragma unroll
          for (int nb = 0; nb < NB; ++nb) {
            acc[nb][0] = fmaf(sa, part[nb][0], acc[nb][0]);
            acc[nb][1] = fmaf(sa, part[nb][1], acc[nb][1]);
            acc[nb][2] = fmaf(sb, part[nb][2], acc[nb][2]);
            acc[nb][3] = fmaf(sb, part[nb][3], acc[nb][3]);
          }
        }
      }
    }
    // ---- reduce over the 8 warps (k-slices) through shared memory, double-buffered: one barrier per tile ----
    float* rbuf = red + (it & 1) * (kGemvWarps * 16 * COLS) + warp * (16 * COLS);
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
      *reinterpret_cast<float2*>(rbuf + g * COLS + 8 * nb + 2 * t) = make_float2(acc[nb][0], acc[nb][1]);
      *reinterpret_cast<float2*>(rbuf + (g + 8) * COLS + 8 * nb + 2 * t) = make_float2(acc[nb][2], acc[nb][3]);
    }
    __syncthreads();
    constexpr int RT = 8 * COLS;  // reducer threads: (r < 8, c), each owns tile rows r and r+8
    if (tid < RT) {
      const int r = tid / COLS, c = tid % COLS;
      const float* rb = red + (it & 1) * (kGemvWarps * 16 * COLS);
      float vA = 0.f, vB = 0.f;
#pragma unroll
      for (int w = 0; w < kGemvWarps; ++w) {
        vA += rb[w * 16 * COLS + r * COLS + c];
        vB += rb[w * 16 * COLS + (r + 8) * COLS + c];
      }
      bool do_epilogue = true;
      if (S > 1) {
        float* part = p.ws_part + ((int64_t)tl * S + s) * (16 * COLS);
        __stcg(part + r * COLS + c, vA);
        __stcg(part + (r + 8) * COLS + c, vB);
        __threadfence();
        asm volatile("bar.sync 1, %0;" ::"n"(RT));
        if (tid == 0) {
          unsigned int old = atomicAdd(p.ws_cnt + tl, 1u);
          s_flag[it & 1] = (old == (unsigned int)(S - 1));
          if (old == (unsigned int)(S - 1)) p.ws_cnt[tl] = 0u;  // self-cleaning for the next launch
        }
        asm volatile("bar.sync 1, %0;" ::"n"(RT));
        do_epilogue = s_flag[it & 1] != 0u;
        if (do_epilogue) {
          __threadfence();
          vA = vB = 0.f;
          for (int ss = 0; ss < S; ++ss) {  // fixed order: deterministic
            const float* q = p.ws_part + ((int64_t)tl * S + ss) * (16 * COLS);
            vA += __ldcg(q + r * COLS + c);
            vB += __ldcg(q + (r + 8) * COLS + c);
          }
        }
      }
      if (do_epilogue && c < B) {
        const int64_t n0 = (int64_t)tl * p.rows_per_tile;
        if (p.rows_per_tile == 8) {
          if (n0 + r < p.N) epilogue_store(p, n0 + r, c, vA, vB);
        } else {
          if (n0 + r < p.N) epilogue_store(p, n0 + r, c, vA, 0.f);
          if (n0 + r + 8 < p.N) epilogue_store(p, n0 + r + 8, c, vB, 0.f);
        }
      }
    }
  };

  int it = 0;
  for (; tile < p.ntiles; tile += 2 * stride) {
    process(bufA, tile, it++);
    load_tile<KS>(bufA, p, tile + 2 * stride, g, t, kb0, cnt);
    if (tile + stride < p.ntiles) process(bufB, tile + stride, it++);
    load_tile<KS>(bufB, p, tile + 3 * stride, g, t, kb0, cnt);
  }
}

// ------------------------------------------------------------------------------------------------------------
// Host side
// ------------------------------------------------------------------------------------------------------------
static std::atomic<int> g_tune_ks{0}, g_tune_occ{0}, g_tune_generic{0};

constexpr size_t kWsCounters = 64 * 1024;           // 16384 tile counters
constexpr size_t kWsXin = 2 * 1024 * 1024;          // staged activations for the generic path: 16 x 65536 bf16
constexpr size_t kWsPartials = 14 * 1024 * 1024;    // split-K partial sums
constexpr size_t kWsTotal = kWsCounters + kWsXin + kWsPartials;

template <int KS, int NB, bool INT4>
static cudaError_t launch_fast(const GemvParams& p, int gx, int S, cudaStream_t st) {
  constexpr int COLS = 8 * NB;
  const size_t smem = kLutBytes + sizeof(float) * (2 * kGemvWarps * 16 * COLS + kGemvWarps * 16 + 16) + 16;
  static PerDeviceOnce once;  // per instantiation
  if (once.need()) {
    cudaError_t e = cudaFuncSetAttribute(qgemv_fast_kernel<KS, NB, INT4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    once.done();
  }
  return launch(qgemv_fast_kernel<KS, NB, INT4>, dim3(gx, S), dim3(kGemvThreads), smem, st, p);
}

template <int NB, bool INT4>
static cudaError_t dispatch_ks(int KS, const GemvParams& p, int gx, int S, cudaStream_t st) {
  switch (KS) {
    case 1: return launch_fast<1, NB, INT4>(p, gx, S, st);
    case 2: return launch_fast<2, NB, INT4>(p, gx, S, st);
    default: return launch_fast<4, NB, INT4>(p, gx, S, st);
  }
}

}  // namespace b200

using namespace b200;

extern "C" {

size_t b200_qgemv_workspace_bytes(void) { return kWsTotal; }

int b200_qgemv_tune(int ksteps, int ctas_per_sm, int force_generic) {
  g_tune_ks = ksteps;
  g_tune_occ = ctas_per_sm;
  g_tune_generic = force_generic;
  return B200_OK;
}

int b200_rmsnorm(const void* x, const void* delta, const void* weight, float eps, void* y, void* sum_out, int64_t rows,
                 int64_t dim, b200_stream_t stream);

int b200_qgemv(const b200_qgemv_args* a, b200_stream_t stream) {
  B200_CHECK_ARG(a != nullptr, "b200_qgemv: null args");
  B200_CHECK_ARG(a->packed && a->scale && a->x && a->y, "b200_qgemv: null packed/scale/x/y pointer");
  B200_CHECK_ARG(a->N > 0 && a->K > 0, "b200_qgemv: N and K must be positive (N=%lld K=%lld)", (long long)a->N, (long long)a->K);
  B200_CHECK_ARG(a->batch >= 1 && a->batch <= 16, "b200_qgemv: batch must be in [1,16] (got %d); use b200_qgemm for prefill", a->batch);
  B200_CHECK_ARG(a->qtype == B200_Q_NF4 || a->qtype == B200_Q_FP4 || a->qtype == B200_Q_INT4, "b200_qgemv: bad qtype %d", a->qtype);
  B200_CHECK_ARG(a->epilogue >= B200_EPI_STORE && a->epilogue <= B200_EPI_SWIGLU, "b200_qgemv: bad epilogue %d", a->epilogue);
  B200_CHECK_ARG(a->epilogue != B200_EPI_SWIGLU || (a->packed2 && a->scale2), "b200_qgemv: SWIGLU needs packed2/scale2");
  B200_CHECK_ARG(a->epilogue != B200_EPI_RESIDUAL || a->residual, "b200_qgemv: RESIDUAL epilogue needs residual");
  B200_CHECK_ARG(a->workspace && a->workspace_bytes >= kWsTotal, "b200_qgemv: workspace too small (%zu < %zu)", a->workspace_bytes, kWsTotal);
  B200_CHECK_ARG(a->x_out == nullptr || a->x_out != a->x, "b200_qgemv: x_out must not alias x");
  if (a->qtype == B200_Q_INT4)
    B200_CHECK_ARG(a->blocksize == 128 || a->blocksize == 256, "b200_qgemv: int4 group must be 128 or 256 (got %d)", a->blocksize);
  else
    B200_CHECK_ARG(a->blocksize >= 2 && a->blocksize % 2 == 0, "b200_qgemv: blocksize must be even (got %d)", a->blocksize);
  cudaStream_t st = (cudaStream_t)stream;
  uint8_t* ws = reinterpret_cast<uint8_t*>(a->workspace);

  const bool aligned = ((reinterpret_cast<uintptr_t>(a->x) | reinterpret_cast<uintptr_t>(a->packed) |
                         reinterpret_cast<uintptr_t>(a->delta) | reinterpret_cast<uintptr_t>(a->x_out) |
                         reinterpret_cast<uintptr_t>(a->norm_weight) | reinterpret_cast<uintptr_t>(a->packed2)) & 15) == 0;
  const bool fast_fmt = (a->qtype == B200_Q_INT4) ? (a->K % a->blocksize == 0) : (a->blocksize == 64);
  const bool fast = fast_fmt && (a->K % 64 == 0) && aligned && !g_tune_generic.load() && a->N >= 8 &&
                    (a->epilogue != B200_EPI_SWIGLU || a->bias == nullptr);

  if (!fast) {
    // generic path: prologue by the standalone norm kernel into the workspace, then one warp per output row
    const __nv_bfloat16* xin = reinterpret_cast<const __nv_bfloat16*>(a->x);
    if (a->delta || a->norm_weight) {
      B200_CHECK_ARG((size_t)a->batch * a->K * 2 <= kWsXin, "b200_qgemv: K too large for the generic path (%lld)", (long long)a->K);
      B200_CHECK_ARG(a->norm_weight != nullptr, "b200_qgemv: generic path needs norm_weight when delta is given");
      __nv_bfloat16* stage = reinterpret_cast<__nv_bfloat16*>(ws + kWsCounters);
      int rc = b200_rmsnorm(a->x, a->delta, a->norm_weight, a->norm_eps, stage, a->x_out, a->batch, a->K, stream);
      if (rc != B200_OK) return rc;
      xin = stage;
    }
    GenericParams g{};
    g.w[0] = reinterpret_cast<const uint8_t*>(a->packed);
    g.s[0] = reinterpret_cast<const uint32_t*>(a->scale);
    g.w[1] = reinterpret_cast<const uint8_t*>(a->packed2);
    g.s[1] = reinterpret_cast<const uint32_t*>(a->scale2);
    g.N = a->N; g.K = a->K; g.blocksize = a->blocksize; g.qtype = a->qtype; g.batch = a->batch; g.epilogue = a->epilogue;
    g.x = xin;
    g.bias = reinterpret_cast<const __nv_bfloat16*>(a->bias);
    g.bias2 = nullptr;
    g.residual = reinterpret_cast<const __nv_bfloat16*>(a->residual);
    g.y = reinterpret_cast<__nv_bfloat16*>(a->y);
    int64_t blocks = (a->N + 7) / 8;
    int64_t cap = (int64_t)sm_count() * 8;
    B200_CUDA(launch(qgemv_generic_kernel, dim3((unsigned)(blocks < cap ? blocks : cap)), dim3(256), 0, st, g));
    return B200_OK;
  }

  GemvParams p{};
  const bool swiglu = a->epilogue == B200_EPI_SWIGLU;
  p.N = a->N; p.K = a->K; p.nkb = (int)(a->K / 64);
  p.qtype = a->qtype;
  if (a->qtype == B200_Q_INT4) { p.sc_per_row = (int)(a->K / a->blocksize); p.sc_shift = a->blocksize == 128 ? 1 : 2; }
  else { p.sc_per_row = p.nkb; p.sc_shift = 0; }
  B200_CHECK_ARG(a->qtype != B200_Q_INT4 || a->K % a->blocksize == 0, "b200_qgemv: int4 fast path needs K %% group == 0");
  p.rows_per_tile = swiglu ? 8 : 16;
  p.ntiles = (int)((a->N + p.rows_per_tile - 1) / p.rows_per_tile);
  p.wA = reinterpret_cast<const uint8_t*>(a->packed);
  p.sA = reinterpret_cast<const uint32_t*>(a->scale);
  if (swiglu) {
    p.wB = reinterpret_cast<const uint8_t*>(a->packed2);
    p.sB = reinterpret_cast<const uint32_t*>(a->scale2);
  } else {
    p.wB = p.wA;  // rows n0+8+g are addressed by row index inside load_tile
    p.sB = p.sA;
  }
  p.x = reinterpret_cast<const __nv_bfloat16*>(a->x);
  p.delta = reinterpret_cast<const __nv_bfloat16*>(a->delta);
  p.x_out = reinterpret_cast<__nv_bfloat16*>(a->x_out);
  p.norm_w = reinterpret_cast<const __nv_bfloat16*>(a->norm_weight);
  p.eps = a->norm_eps;
  p.batch = a->batch;
  p.epilogue = a->epilogue;
  p.bias = reinterpret_cast<const __nv_bfloat16*>(a->bias);
  p.bias2 = nullptr;
  p.residual = reinterpret_cast<const __nv_bfloat16*>(a->residual);
  p.y = reinterpret_cast<__nv_bfloat16*>(a->y);
  p.ws_cnt = reinterpret_cast<unsigned int*>(ws);
  p.ws_part = reinterpret_cast<float*>(ws + kWsCounters + kWsXin);

  const int NB = a->batch > 8 ? 2 : 1;
  int KS = g_tune_ks.load();
  if (KS != 1 && KS != 2 && KS != 4) KS = p.nkb <= 8 ? 1 : (p.nkb <= 16 ? 2 : 4);
  if (NB == 2 && KS == 4) KS = 2;  // register budget: 2 column groups x 2 k-blocks x 8 fragment registers
  const int S = (p.nkb + kGemvWarps * KS - 1) / (kGemvWarps * KS);
  int occ = g_tune_occ.load();
  if (occ < 1 || occ > 4) occ = 2;
  int gx = (sm_count() * occ) / S;
  if (gx < 1) gx = 1;
  if (gx > p.ntiles) gx = p.ntiles;
  B200_CHECK_ARG((size_t)p.ntiles * sizeof(unsigned int) <= kWsCounters, "b200_qgemv: N too large (%lld rows)", (long long)a->N);
  B200_CHECK_ARG(S == 1 || (size_t)p.ntiles * S * 16 * 8 * NB * sizeof(float) <= kWsPartials,
                 "b200_qgemv: split-K workspace too small for N=%lld K=%lld", (long long)a->N, (long long)a->K);
  cudaError_t e;
  if (a->qtype == B200_Q_INT4)
    e = NB == 2 ? dispatch_ks<2, true>(KS, p, gx, S, st) : dispatch_ks<1, true>(KS, p, gx, S, st);
  else
    e = NB == 2 ? dispatch_ks<2, false>(KS, p, gx, S, st) : dispatch_ks<1, false>(KS, p, gx, S, st);
  if (e != cudaSuccess) {
    set_error("b200_qgemv: launch failed: %s", cudaGetErrorString(e));
    return B200_ERR_CUDA;
  }
  return B200_OK;
}

}  // extern "C"
