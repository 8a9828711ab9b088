// Streaming fused dequant-GEMV over the TILED weight layout (csrc/tile.cu), batch 1..16.
//
// Replaces bnb.nn.Linear4bit.forward at lit_gpt/model.py:619,656,713,714,716,519.
//
// One CTA (8 warps) owns whole output rows: a "pair" is two 16-row tiles (rows 32p..32p+31, or, for SwiGLU, tile p
// of the gate matrix and tile p of the up matrix), and the 8 warps split K between them, so there is NO split-K
// across CTAs, no global workspace, no atomics and no fences.  Per 64-wide k-block and tile a warp issues ONE
// fully coalesced 512-byte load (ld.global.nc, L1 no-allocate, L2 evict-first) plus one 64-byte scale load, kept
// DEPTH items ahead in registers; the first DEPTH items are requested before griddepcontrol.wait, so under
// programmatic dependent launch the weight stream of kernel N+1 is already in flight while kernel N drains.
// Each CTA also issues cp.async.bulk.prefetch.L2 for a caller-given range (the NEXT kernel's weights), which keeps
// HBM busy across the kernel boundary.
// Dequant: one PRMT builds a shared-memory address from a packed byte, one conflict-free LDS from a per-lane
// replicated 256-entry byte -> bf16x2 table returns two code values in mma.sync A-fragment order; mma.sync
// m16n8k16 (bf16 x bf16 -> fp32) does the MACs against activation B-fragments read from shared memory (shared by
// the two tiles of a pair); the per-64 absmax (or int4 {scale, zero}) is applied in fp32 per k-block.
// Prologue: residual add + RMSNorm of the activation (lit_gpt/rmsnorm.py:19-25, model.py:591) into shared memory.
// Epilogue: bias / residual add / SwiGLU with the reference's bf16 rounding points (lit_gpt/model.py:712-716).
#include "common.cuh"

namespace b200 {

struct StreamParams {
  const uint8_t* tA;   // tiled matrix (tile.cu layout)
  const uint8_t* tB;   // second tiled matrix for SwiGLU (else == tA)
  int64_t N, K;
  int nkb;             // K / 64
  int ntiles;          // ceil(N / 16)
  int npairs;          // SwiGLU: ntiles, else ceil(ntiles / 2)
  int swiglu;
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
  int xstride;         // shared-memory row stride of the activation, in bf16 elements (K + 8)
  const uint8_t* pf;   // L2 prefetch range (next kernel's weights), may be null
  int64_t pf_bytes;
};

constexpr int kSThreads = 256;
constexpr int kSWarps = 8;
constexpr int kSLutBytes = 256 * 256;  // entry e, lane l at byte e*256 + l*4

__device__ __forceinline__ void mma16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

__device__ __forceinline__ uint64_t make_evict_first_policy() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ uint4 ldg_w_v4(const void* p, uint64_t pol) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p), "l"(pol));
  return r;
}
__device__ __forceinline__ uint2 ldg_w_v2(const void* p, uint64_t pol) {
  uint2 r;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.u32 {%0,%1}, [%2], %3;" : "=r"(r.x), "=r"(r.y) : "l"(p), "l"(pol));
  return r;
}
__device__ __forceinline__ void prefetch_l2_bulk(const void* p, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}

__device__ __forceinline__ float silu_bf16_s(float a) { return round_bf16(a / (1.0f + expf(-a))); }

struct SItem {
  uint4 w0, w1;  // 16 B of tile 0 / tile 1: {row g bytes 0..7, row g+8 bytes 0..7}
  uint2 s0, s1;  // scale words of rows g, g+8 for tile 0 / tile 1
};

// NB: batch column groups of 8 (batch <= 8*NB); INT4: int4 group-128 records; D: load-ahead depth in items.
template <int NB, bool INT4, int D>
__global__ void __launch_bounds__(kSThreads, 2) qgemv_stream_kernel(const StreamParams p) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  constexpr int COLS = 8 * NB;
  constexpr int RED_ONE = kSWarps * 2 * 16 * COLS;  // floats per reduction buffer
  uint32_t* lut = reinterpret_cast<uint32_t*>(smem_raw);
  float* red = reinterpret_cast<float*>(smem_raw + kSLutBytes);                 // [2][8 warps][2 tiles][16][COLS]
  float* s_ssq = red + 2 * RED_ONE;                                             // [8 warps][16]
  float* s_rstd = s_ssq + kSWarps * 16;                                         // [16]
  float* xsum_s = s_rstd + 16;                                                  // INT4: [batch][nkb]
  __nv_bfloat16* x_s = reinterpret_cast<__nv_bfloat16*>(xsum_s + (INT4 ? ((p.batch * p.nkb + 3) & ~3) : 0));  // [batch][xstride]

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int B = p.batch;
  const uint64_t pol = make_evict_first_policy();

  // this warp's k-blocks [kb0, kb1)
  int kb0, kb1;
  if (INT4) { kb0 = 2 * ((warp * (p.nkb >> 1)) / kSWarps); kb1 = 2 * (((warp + 1) * (p.nkb >> 1)) / kSWarps); }
  else { kb0 = (warp * p.nkb) / kSWarps; kb1 = ((warp + 1) * p.nkb) / kSWarps; }
  const int len = kb1 - kb0;
  constexpr int REC = INT4 ? 1088 : 576;
  const int nrec = INT4 ? (p.nkb >> 1) : p.nkb;

  auto load_item = [&](SItem& it, int pair, int kb) {
    const int tile0 = p.swiglu ? pair : 2 * pair;
    const int tile1 = p.swiglu ? pair : 2 * pair + 1;
    const int rec = INT4 ? (kb >> 1) : kb;
    const int woff = (INT4 ? (kb & 1) * 512 : 0) + lane * 16;
    const int soff = (INT4 ? 1024 : 512) + g * 8;
    const uint8_t* r0 = p.tA + ((int64_t)tile0 * nrec + rec) * REC;
    it.w0 = ldg_w_v4(r0 + woff, pol);
    if (!INT4 || !(kb & 1)) it.s0 = ldg_w_v2(r0 + soff, pol);
    if (tile1 < p.ntiles) {
      const uint8_t* r1 = p.tB + ((int64_t)tile1 * nrec + rec) * REC;
      it.w1 = ldg_w_v4(r1 + woff, pol);
      if (!INT4 || !(kb & 1)) it.s1 = ldg_w_v2(r1 + soff, pol);
    } else {
      it.w1 = make_uint4(0u, 0u, 0u, 0u);
      it.s1 = make_uint2(0u, 0u);
    }
  };

  // ---- input-independent prologue: first D items in flight, L2 prefetch of the next kernel's weights, LUT ----
  SItem ring[D];
#pragma unroll
  for (int d = 0; d < D; ++d) { ring[d].w0 = ring[d].w1 = make_uint4(0u, 0u, 0u, 0u); ring[d].s0 = ring[d].s1 = make_uint2(0u, 0u); }
  int ld_pair = blockIdx.x, ld_kb = kb0;
  auto ld_advance = [&]() { if (++ld_kb >= kb1) { ld_kb = kb0; ld_pair += gridDim.x; } };
  if (len > 0) {
#pragma unroll
    for (int d = 0; d < D; ++d) {
      if (ld_pair < p.npairs) { load_item(ring[d], ld_pair, ld_kb); ld_advance(); }
    }
  }
  if (p.pf != nullptr && lane == 0) {
    // CTA share, split over the 8 warps, in pieces of <= 8 KB (size and address multiples of 16)
    const int64_t per_cta = (((p.pf_bytes + gridDim.x - 1) / gridDim.x) + 1023) & ~(int64_t)1023;
    const int64_t per_warp = per_cta / kSWarps;  // multiple of 128
    int64_t beg = (int64_t)blockIdx.x * per_cta + (int64_t)warp * per_warp;
    int64_t end = beg + per_warp;
    if (end > p.pf_bytes) end = p.pf_bytes & ~(int64_t)15;
    for (; beg < end; beg += 8192) {
      const int64_t n = end - beg < 8192 ? end - beg : 8192;
      prefetch_l2_bulk(p.pf + beg, (uint32_t)n);
    }
  }
  for (int i = tid; i < 256 * 32; i += kSThreads) {
    const int e = i >> 5, l = i & 31;
    float lo, hi;
    if (INT4) { lo = (float)(e & 15); hi = (float)(e >> 4); }                      // k even = LOW nibble
    else { lo = code_value(p.qtype, e >> 4); hi = code_value(p.qtype, e & 15); }   // even element = HIGH nibble
    lut[e * 64 + l] = pack_bf16(lo, hi);
  }
  griddep_launch();
  griddep_wait();

  // ---- activation prologue: (x + delta) -> optional RMSNorm -> bf16 rows in shared memory ---------------------
  const bool designated = blockIdx.x == 0;
  const int chunks = (int)(p.K >> 3);
  for (int b = 0; b < B; ++b) {
    float ss = 0.f;
    for (int c = tid; c < chunks; c += kSThreads) {
      uint4 v = *reinterpret_cast<const uint4*>(p.x + (int64_t)b * p.K + c * 8);
      if (p.delta) {
        const uint4 dv = *reinterpret_cast<const uint4*>(p.delta + (int64_t)b * p.K + c * 8);
        v.x = pack_bf16(bf16lo(v.x) + bf16lo(dv.x), bf16hi(v.x) + bf16hi(dv.x));
        v.y = pack_bf16(bf16lo(v.y) + bf16lo(dv.y), bf16hi(v.y) + bf16hi(dv.y));
        v.z = pack_bf16(bf16lo(v.z) + bf16lo(dv.z), bf16hi(v.z) + bf16hi(dv.z));
        v.w = pack_bf16(bf16lo(v.w) + bf16lo(dv.w), bf16hi(v.w) + bf16hi(dv.w));
        if (p.x_out && designated) *reinterpret_cast<uint4*>(p.x_out + (int64_t)b * p.K + c * 8) = v;
      }
      *reinterpret_cast<uint4*>(x_s + b * p.xstride + c * 8) = v;
      const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) { const float a = bf16lo(w4[i]), c2 = bf16hi(w4[i]); ss = fmaf(a, a, ss); ss = fmaf(c2, c2, ss); }
    }
    if (p.norm_w) {
      ss = warp_sum(ss);
      if (lane == 0) s_ssq[warp * 16 + b] = ss;
    }
  }
  if (p.norm_w) {
    __syncthreads();
    if (tid < B) {
      float tot = 0.f;
#pragma unroll
      for (int w = 0; w < kSWarps; ++w) tot += s_ssq[w * 16 + tid];
      s_rstd[tid] = rsqrtf(tot / (float)p.K + p.eps);
    }
    __syncthreads();
    for (int b = 0; b < B; ++b) {
      const float rstd = s_rstd[b];
      for (int c = tid; c < chunks; c += kSThreads) {  // same thread owns the same chunk as above
        uint4 v = *reinterpret_cast<uint4*>(x_s + b * p.xstride + c * 8);
        const uint4 nw = *reinterpret_cast<const uint4*>(p.norm_w + c * 8);
        // weight * (x * rstd), fp32, one cast (rmsnorm.py:24-25)
        v.x = pack_bf16(bf16lo(nw.x) * (bf16lo(v.x) * rstd), bf16hi(nw.x) * (bf16hi(v.x) * rstd));
        v.y = pack_bf16(bf16lo(nw.y) * (bf16lo(v.y) * rstd), bf16hi(nw.y) * (bf16hi(v.y) * rstd));
        v.z = pack_bf16(bf16lo(nw.z) * (bf16lo(v.z) * rstd), bf16hi(nw.z) * (bf16hi(v.z) * rstd));
        v.w = pack_bf16(bf16lo(nw.w) * (bf16lo(v.w) * rstd), bf16hi(nw.w) * (bf16hi(v.w) * rstd));
        *reinterpret_cast<uint4*>(x_s + b * p.xstride + c * 8) = v;
      }
    }
  }
  __syncthreads();  // x_s and the LUT are published
  if (INT4) {       // per (batch row, k-block) activation sums for the zero-point term
    for (int i = tid; i < B * p.nkb; i += kSThreads) {
      const int b = i / p.nkb, kb = i % p.nkb;
      const uint32_t* px = reinterpret_cast<const uint32_t*>(x_s + b * p.xstride + kb * 64);
      float s = 0.f;
#pragma unroll 8
      for (int j = 0; j < 32; ++j) s += bf16lo(px[j]) + bf16hi(px[j]);
      xsum_s[i] = s;
    }
    __syncthreads();
  }

  // ---- main loop ------------------------------------------------------------------------------------------------
  const uint32_t lane_off = (uint32_t)lane << 2;
  const uint8_t* lut_b = reinterpret_cast<const uint8_t*>(lut);
  auto LUT = [&](uint32_t word, uint32_t sel) -> uint32_t {
    // address = (byte_i << 8) | (lane << 2): PRMT picks lane_off.b0, word.b_i, 0, 0
    return *reinterpret_cast<const uint32_t*>(lut_b + __byte_perm(word, lane_off, sel));
  };

  float acc0[NB][4], acc1[NB][4];
#pragma unroll
  for (int nb = 0; nb < NB; ++nb) {
#pragma unroll
    for (int i = 0; i < 4; ++i) acc0[nb][i] = acc1[nb][i] = 0.f;
  }
  uint2 hold0 = make_uint2(0u, 0u), hold1 = make_uint2(0u, 0u);  // INT4: {scale, zero} of the current group

  auto compute = [&](const SItem& it, int kb) {
    uint32_t xb[NB][8];
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
      const int b = g + 8 * nb;
      if (b < B) {
        const uint4* px = reinterpret_cast<const uint4*>(x_s + b * p.xstride + kb * 64 + t * 16);
        const uint4 v0 = px[0], v1 = px[1];
        xb[nb][0] = v0.x; xb[nb][1] = v0.y; xb[nb][2] = v0.z; xb[nb][3] = v0.w;
        xb[nb][4] = v1.x; xb[nb][5] = v1.y; xb[nb][6] = v1.z; xb[nb][7] = v1.w;
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) xb[nb][i] = 0u;
      }
    }
    float p0[NB][4], p1[NB][4];
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
#pragma unroll
      for (int i = 0; i < 4; ++i) p0[nb][i] = p1[nb][i] = 0.f;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t sel0 = (j & 1) ? 0x5524u : 0x5504u, sel1 = (j & 1) ? 0x5534u : 0x5514u;
      {
        const uint32_t wa = (j < 2) ? it.w0.x : it.w0.y, wb = (j < 2) ? it.w0.z : it.w0.w;
        const uint32_t a0 = LUT(wa, sel0), a2 = LUT(wa, sel1), a1 = LUT(wb, sel0), a3 = LUT(wb, sel1);
#pragma unroll
        for (int nb = 0; nb < NB; ++nb) mma16816(p0[nb], a0, a1, a2, a3, xb[nb][2 * j], xb[nb][2 * j + 1]);
      }
      {
        const uint32_t wa = (j < 2) ? it.w1.x : it.w1.y, wb = (j < 2) ? it.w1.z : it.w1.w;
        const uint32_t a0 = LUT(wa, sel0), a2 = LUT(wa, sel1), a1 = LUT(wb, sel0), a3 = LUT(wb, sel1);
#pragma unroll
        for (int nb = 0; nb < NB; ++nb) mma16816(p1[nb], a0, a1, a2, a3, xb[nb][2 * j], xb[nb][2 * j + 1]);
      }
    }
    if (INT4) {
      if (!(kb & 1)) { hold0 = it.s0; hold1 = it.s1; }
      const __half2 h0a = *reinterpret_cast<const __half2*>(&hold0.x), h0b = *reinterpret_cast<const __half2*>(&hold0.y);
      const __half2 h1a = *reinterpret_cast<const __half2*>(&hold1.x), h1b = *reinterpret_cast<const __half2*>(&hold1.y);
      const float s0a = __low2float(h0a), z0a = __high2float(h0a), s0b = __low2float(h0b), z0b = __high2float(h0b);
      const float s1a = __low2float(h1a), z1a = __high2float(h1a), s1b = __low2float(h1b), z1b = __high2float(h1b);
#pragma unroll
      for (int nb = 0; nb < NB; ++nb) {
        const int c0 = 2 * t + 8 * nb;
        const float xs0 = c0 < B ? xsum_s[c0 * p.nkb + kb] : 0.f;
        const float xs1 = c0 + 1 < B ? xsum_s[(c0 + 1) * p.nkb + kb] : 0.f;
        acc0[nb][0] = fmaf(s0a, p0[nb][0] - z0a * xs0, acc0[nb][0]);
        acc0[nb][1] = fmaf(s0a, p0[nb][1] - z0a * xs1, acc0[nb][1]);
        acc0[nb][2] = fmaf(s0b, p0[nb][2] - z0b * xs0, acc0[nb][2]);
        acc0[nb][3] = fmaf(s0b, p0[nb][3] - z0b * xs1, acc0[nb][3]);
        acc1[nb][0] = fmaf(s1a, p1[nb][0] - z1a * xs0, acc1[nb][0]);
        acc1[nb][1] = fmaf(s1a, p1[nb][1] - z1a * xs1, acc1[nb][1]);
        acc1[nb][2] = fmaf(s1b, p1[nb][2] - z1b * xs0, acc1[nb][2]);
        acc1[nb][3] = fmaf(s1b, p1[nb][3] - z1b * xs1, acc1[nb][3]);
      }
    } else {
      const float s0a = __uint_as_float(it.s0.x), s0b = __uint_as_float(it.s0.y);
      const float s1a = __uint_as_float(it.s1.x), s1b = __uint_as_float(it.s1.y);
#pragma unroll
      for (int nb = 0; nb < NB; ++nb) {
        acc0[nb][0] = fmaf(s0a, p0[nb][0], acc0[nb][0]);
        acc0[nb][1] = fmaf(s0a, p0[nb][1], acc0[nb][1]);
        acc0[nb][2] = fmaf(s0b, p0[nb][2], acc0[nb][2]);
        acc0[nb][3] = fmaf(s0b, p0[nb][3], acc0[nb][3]);
        acc1[nb][0] = fmaf(s1a, p1[nb][0], acc1[nb][0]);
        acc1[nb][1] = fmaf(s1a, p1[nb][1], acc1[nb][1]);
        acc1[nb][2] = fmaf(s1b, p1[nb][2], acc1[nb][2]);
        acc1[nb][3] = fmaf(s1b, p1[nb][3], acc1[nb][3]);
      }
    }
  };

  auto store_out = [&](int64_t n, int c, float v) {  // STORE / RESIDUAL epilogues for one (row, batch column)
    float h = v + (p.bias ? __bfloat162float(p.bias[n]) : 0.f);
    if (p.epilogue == B200_EPI_RESIDUAL) h = round_bf16(h) + __bfloat162float(p.residual[(int64_t)c * p.N + n]);
    p.y[(int64_t)c * p.N + n] = __float2bfloat16_rn(h);
  };

  int slot = 0, it_no = 0;
  for (int pair = blockIdx.x; pair < p.npairs; pair += gridDim.x, ++it_no) {
    int kb = kb0;
    while (kb < kb1) {
#pragma unroll
      for (int d = 0; d < D; ++d) {
        if (d == slot && kb < kb1) {  // warp-uniform
          compute(ring[d], kb);
          if (ld_pair < p.npairs) { load_item(ring[d], ld_pair, ld_kb); ld_advance(); }
          ++kb;
          slot = (d + 1 == D) ? 0 : d + 1;
        }
      }
    }
    // ---- reduce the 8 k-slices through shared memory (double-buffered: one barrier per pair) ------------------
    float* rbuf = red + (it_no & 1) * RED_ONE + warp * (2 * 16 * COLS);
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
      *reinterpret_cast<float2*>(rbuf + g * COLS + 8 * nb + 2 * t) = make_float2(acc0[nb][0], acc0[nb][1]);
      *reinterpret_cast<float2*>(rbuf + (g + 8) * COLS + 8 * nb + 2 * t) = make_float2(acc0[nb][2], acc0[nb][3]);
      *reinterpret_cast<float2*>(rbuf + 16 * COLS + g * COLS + 8 * nb + 2 * t) = make_float2(acc1[nb][0], acc1[nb][1]);
      *reinterpret_cast<float2*>(rbuf + 16 * COLS + (g + 8) * COLS + 8 * nb + 2 * t) = make_float2(acc1[nb][2], acc1[nb][3]);
#pragma unroll
      for (int i = 0; i < 4; ++i) acc0[nb][i] = acc1[nb][i] = 0.f;
    }
    __syncthreads();
    if (tid < 16 * COLS) {
      const int r = tid / COLS, c = tid % COLS;
      const float* rb = red + (it_no & 1) * RED_ONE;
      float v0 = 0.f, v1 = 0.f;
#pragma unroll
      for (int w = 0; w < kSWarps; ++w) {
        v0 += rb[w * (2 * 16 * COLS) + r * COLS + c];
        v1 += rb[w * (2 * 16 * COLS) + 16 * COLS + r * COLS + c];
      }
      if (c < B) {
        if (p.swiglu) {
          const int64_t n = (int64_t)pair * 16 + r;
          if (n < p.N) {
            const float a = round_bf16(v0 + (p.bias ? __bfloat162float(p.bias[n]) : 0.f));
            const float b = round_bf16(v1 + (p.bias2 ? __bfloat162float(p.bias2[n]) : 0.f));
            p.y[(int64_t)c * p.N + n] = __float2bfloat16_rn(silu_bf16_s(a) * b);
          }
        } else {
          const int64_t n = (int64_t)pair * 32 + r;
          if (n < p.N) store_out(n, c, v0);
          if (n + 16 < p.N) store_out(n + 16, c, v1);
        }
      }
    }
  }
}

static std::atomic<int> g_stream_depth{0}, g_stream_occ{0}, g_stream_pf{1};

template <int NB, bool INT4, int D>
static cudaError_t launch_stream(const StreamParams& p, int gx, size_t smem, cudaStream_t st) {
  static size_t configured = 0;  // per instantiation: largest dynamic smem size opted into so far
  if (smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(qgemv_stream_kernel<NB, INT4, D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024));
    if (e != cudaSuccess) return e;
    configured = 227 * 1024;
  }
  return launch(qgemv_stream_kernel<NB, INT4, D>, dim3(gx), dim3(kSThreads), smem, st, p);
}

template <int NB, bool INT4>
static cudaError_t dispatch_depth(int D, const StreamParams& p, int gx, size_t smem, cudaStream_t st) {
  if (NB == 1 && D >= 6) return launch_stream<NB, INT4, 6>(p, gx, smem, st);
  if (D <= 2) return launch_stream<NB, INT4, 2>(p, gx, smem, st);
  return launch_stream<NB, INT4, 4>(p, gx, smem, st);
}

}  // namespace b200

using namespace b200;

extern "C" {

int b200_qgemv_tiled_tune(int depth, int ctas_per_sm, int l2_prefetch) {
  g_stream_depth = depth;
  g_stream_occ = ctas_per_sm;
  g_stream_pf = l2_prefetch;
  return B200_OK;
}

int b200_qgemv_tiled(const b200_qgemv_tiled_args* a, b200_stream_t stream) {
  B200_CHECK_ARG(a != nullptr, "b200_qgemv_tiled: null args");
  B200_CHECK_ARG(a->tiled && a->x && a->y, "b200_qgemv_tiled: null tiled/x/y pointer");
  B200_CHECK_ARG(a->N > 0 && a->K > 0, "b200_qgemv_tiled: N and K must be positive (N=%lld K=%lld)", (long long)a->N, (long long)a->K);
  B200_CHECK_ARG(a->batch >= 1 && a->batch <= 16, "b200_qgemv_tiled: batch must be in [1,16] (got %d); use b200_qgemm for prefill", a->batch);
  B200_CHECK_ARG(a->qtype == B200_Q_NF4 || a->qtype == B200_Q_FP4 || a->qtype == B200_Q_INT4, "b200_qgemv_tiled: bad qtype %d", a->qtype);
  B200_CHECK_ARG(a->epilogue >= B200_EPI_STORE && a->epilogue <= B200_EPI_SWIGLU, "b200_qgemv_tiled: bad epilogue %d", a->epilogue);
  B200_CHECK_ARG(a->epilogue != B200_EPI_SWIGLU || a->tiled2, "b200_qgemv_tiled: SWIGLU needs tiled2");
  B200_CHECK_ARG(a->epilogue != B200_EPI_RESIDUAL || a->residual, "b200_qgemv_tiled: RESIDUAL epilogue needs residual");
  B200_CHECK_ARG(a->x_out == nullptr || a->x_out != a->x, "b200_qgemv_tiled: x_out must not alias x");
  const bool int4 = a->qtype == B200_Q_INT4;
  B200_CHECK_ARG(a->K % (int4 ? 128 : 64) == 0, "b200_qgemv_tiled: K must be a multiple of %d (got %lld)", int4 ? 128 : 64, (long long)a->K);
  B200_CHECK_ARG(((reinterpret_cast<uintptr_t>(a->x) | reinterpret_cast<uintptr_t>(a->tiled) | reinterpret_cast<uintptr_t>(a->tiled2) |
                   reinterpret_cast<uintptr_t>(a->delta) | reinterpret_cast<uintptr_t>(a->x_out) |
                   reinterpret_cast<uintptr_t>(a->norm_weight) | reinterpret_cast<uintptr_t>(a->prefetch)) & 15) == 0,
                 "b200_qgemv_tiled: pointers must be 16-byte aligned");

  StreamParams p{};
  const bool swiglu = a->epilogue == B200_EPI_SWIGLU;
  p.tA = reinterpret_cast<const uint8_t*>(a->tiled);
  p.tB = swiglu ? reinterpret_cast<const uint8_t*>(a->tiled2) : p.tA;
  p.N = a->N; p.K = a->K; p.nkb = (int)(a->K / 64);
  p.ntiles = (int)((a->N + 15) / 16);
  p.npairs = swiglu ? p.ntiles : (p.ntiles + 1) / 2;
  p.swiglu = swiglu; p.qtype = a->qtype;
  p.x = reinterpret_cast<const __nv_bfloat16*>(a->x);
  p.delta = reinterpret_cast<const __nv_bfloat16*>(a->delta);
  p.x_out = reinterpret_cast<__nv_bfloat16*>(a->x_out);
  p.norm_w = reinterpret_cast<const __nv_bfloat16*>(a->norm_weight);
  p.eps = a->norm_eps; p.batch = a->batch; p.epilogue = a->epilogue;
  p.bias = reinterpret_cast<const __nv_bfloat16*>(a->bias);
  p.bias2 = reinterpret_cast<const __nv_bfloat16*>(a->bias2);
  p.residual = reinterpret_cast<const __nv_bfloat16*>(a->residual);
  p.y = reinterpret_cast<__nv_bfloat16*>(a->y);
  p.xstride = (int)a->K + 8;
  if (g_stream_pf.load() && a->prefetch && a->prefetch_bytes >= 16) { p.pf = reinterpret_cast<const uint8_t*>(a->prefetch); p.pf_bytes = a->prefetch_bytes; }

  const int NB = a->batch > 8 ? 2 : 1;
  const int COLS = 8 * NB;
  size_t smem = kSLutBytes + sizeof(float) * (2 * kSWarps * 2 * 16 * COLS + kSWarps * 16 + 16);
  if (int4) smem += sizeof(float) * (((size_t)a->batch * p.nkb + 3) & ~(size_t)3);
  smem += (size_t)a->batch * p.xstride * 2;
  if (smem > 227 * 1024) {
    set_error("b200_qgemv_tiled: batch %d x K %lld needs %zu bytes of shared memory (> 227 KB)", a->batch, (long long)a->K, smem);
    return B200_ERR_UNSUPPORTED;
  }
  int occ = g_stream_occ.load();
  if (occ < 1 || occ > 2) occ = 1;
  if (smem > 113 * 1024) occ = 1;
  int gx = sm_count() * occ;
  if (gx > p.npairs) gx = p.npairs;
  int D = g_stream_depth.load();
  if (D <= 0) D = 4;
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e;
  if (int4) e = NB == 2 ? dispatch_depth<2, true>(D, p, gx, smem, st) : dispatch_depth<1, true>(D, p, gx, smem, st);
  else e = NB == 2 ? dispatch_depth<2, false>(D, p, gx, smem, st) : dispatch_depth<1, false>(D, p, gx, smem, st);
  if (e != cudaSuccess) {
    set_error("b200_qgemv_tiled: launch failed: %s", cudaGetErrorString(e));
    return B200_ERR_CUDA;
  }
  return B200_OK;
}

}  // extern "C"
