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
  int stages;          // ring stages per warp (2..4)
  int upp;             // work units per pair (NF4/FP4: k-blocks; INT4: 128-wide groups)
  int maxslots;        // stream-K: max groups contributing to one pair
  float* ws_part;      // [npairs][maxslots][2][16][COLS] partial sums of pairs cut by a group boundary
  unsigned int* ws_cnt;  // [npairs] arrival counters (self-cleaning)
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

// ---- mbarrier / bulk-copy (TMA) helpers ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  } while (!done);
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar, uint64_t pol) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar), "l"(pol) : "memory");
}

// Stream-K work split: the (pair, k-unit) space is cut into equal contiguous ranges, one per group of 8 warps, so
// every SM streams the same number of bytes whatever the shape.  A group walks the segments (pair, [a, b)) of its
// range; inside a segment the 8 warps split [a, b) evenly.
struct SegCursor {
  int cur, end;   // group range, in units
  int pair, a, b; // current segment: units [a, b) of `pair`
  int kb0, kb1;   // this warp's k-blocks inside the segment
};
__device__ __forceinline__ void seg_load(SegCursor& s, int upp, int bpu, int warp) {
  s.pair = s.cur / upp;
  s.a = s.cur - s.pair * upp;
  const int n = min(upp - s.a, s.end - s.cur);
  s.b = s.a + n;
  s.kb0 = bpu * (s.a + (n * warp) / 8);
  s.kb1 = bpu * (s.a + (n * (warp + 1)) / 8);
}
__device__ __forceinline__ bool seg_next(SegCursor& s, int upp, int bpu, int warp) {
  s.cur += s.b - s.a;
  if (s.cur >= s.end) return false;
  seg_load(s, upp, bpu, warp);
  return true;
}

constexpr int kGroups = 2;                      // row groups per CTA (each: 8 warps splitting K for one pair)
constexpr int kSThreadsAll = kGroups * kSThreads;
constexpr int kChunkBlocks = 2;                 // k-blocks per ring stage (per tile)
constexpr int kMaxStages = 4;

// Shared-memory layout (dynamic): [LUT 64 KB][mbarriers][red: groups x 8 warps x 2 tiles x 16 x COLS][ssq][rstd][xsum]
// [x rows][ring: 16 warps x stages x stage_bytes].
template <int NB, bool INT4>
__global__ void __launch_bounds__(kSThreadsAll, 1) qgemv_stream_kernel(const StreamParams p) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  constexpr int COLS = 8 * NB;
  constexpr int RED_GROUP = kSWarps * 2 * 16 * COLS;       // floats per group
  constexpr int C = kChunkBlocks;
  constexpr int BLK_BYTES = INT4 ? 544 : 576;              // bytes per k-block per tile inside a record stream
  constexpr int STAGE_TILE = C * BLK_BYTES;                // one tile's part of a stage
  constexpr int STAGE = 2 * STAGE_TILE;                    // two tiles
  uint32_t* lut = reinterpret_cast<uint32_t*>(smem_raw);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + kSLutBytes);               // [16 warps][kMaxStages]
  float* red = reinterpret_cast<float*>(bars + kGroups * kSWarps * kMaxStages);      // [groups][RED_GROUP]
  float* s_ssq = red + kGroups * RED_GROUP;                                          // [16 warps][16]
  float* s_rstd = s_ssq + kGroups * kSWarps * 16;                                    // [16]
  uint32_t* s_flag = reinterpret_cast<uint32_t*>(s_rstd + 16);                       // [groups] (padded to 16)
  float* xsum_s = s_rstd + 32;                                                       // INT4: [batch][nkb]
  __nv_bfloat16* x_s = reinterpret_cast<__nv_bfloat16*>(xsum_s + (INT4 ? ((p.batch * p.nkb + 3) & ~3) : 0));
  uint8_t* ring_base = reinterpret_cast<uint8_t*>(x_s) + (((size_t)p.batch * p.xstride * 2 + 127) & ~(size_t)127);

  const int tid = threadIdx.x, lane = tid & 31, warp_all = tid >> 5;
  const int grp = warp_all / kSWarps, warp = warp_all % kSWarps;   // warp: k-slice index inside the group
  const int gtid = tid - grp * kSThreads;                          // thread index inside the group
  const int g = lane >> 2, t = lane & 3;
  const int B = p.batch;
  const int S = p.stages;
  const uint64_t pol = make_evict_first_policy();

  constexpr int BPU = INT4 ? 2 : 1;                                 // k-blocks per work unit
  const int ngroups = gridDim.x * kGroups, gidx = blockIdx.x * kGroups + grp;
  const int64_t Wtot = (int64_t)p.npairs * p.upp;
  auto bnd = [&](int j) -> int { return (int)((Wtot * j) / ngroups); };
  const int u_begin = bnd(gidx), u_end = bnd(gidx + 1);

  uint8_t* my_ring = ring_base + (size_t)warp_all * S * STAGE;
  const uint32_t my_ring_u32 = smem_u32(my_ring);
  const uint32_t my_bars = smem_u32(bars + warp_all * kMaxStages);

  // producer cursor (lane 0 of every warp feeds its own ring)
  SegCursor ps;
  ps.cur = u_begin; ps.end = u_end; ps.pair = 0; ps.a = ps.b = 0; ps.kb0 = ps.kb1 = 0;
  bool pr_live = u_begin < u_end;
  if (pr_live) seg_load(ps, p.upp, BPU, warp);
  int pr_cb = ps.kb0, pr_stage = 0;
  auto produce = [&]() {  // issue the bulk copies of the next chunk into stage pr_stage (caller guarantees it is free)
    while (pr_live && pr_cb >= ps.kb1) {  // skip exhausted / empty slices
      pr_live = seg_next(ps, p.upp, BPU, warp);
      pr_cb = ps.kb0;
    }
    if (!pr_live) return;
    const int nblk = min(C, ps.kb1 - pr_cb);
    const int tile0 = p.swiglu ? ps.pair : 2 * ps.pair;
    const int tile1 = p.swiglu ? ps.pair : 2 * ps.pair + 1;
    const bool has1 = tile1 < p.ntiles;
    const uint32_t bytes = (uint32_t)nblk * BLK_BYTES;
    const uint32_t bar = my_bars + pr_stage * 8;
    const uint32_t dst = my_ring_u32 + pr_stage * STAGE;
    mbar_expect_tx(bar, has1 ? 2 * bytes : bytes);
    bulk_g2s(dst, p.tA + ((int64_t)tile0 * p.nkb + pr_cb) * BLK_BYTES, bytes, bar, pol);
    if (has1) bulk_g2s(dst + STAGE_TILE, p.tB + ((int64_t)tile1 * p.nkb + pr_cb) * BLK_BYTES, bytes, bar, pol);
    pr_cb += C;
    pr_stage = pr_stage + 1 == S ? 0 : pr_stage + 1;
  };

  // ---- input-independent prologue ------------------------------------------------------------------------------
  if (lane == 0) {
    for (int s = 0; s < kMaxStages; ++s) mbar_init(my_bars + s * 8, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    for (int s = 0; s < S; ++s) produce();
    if (p.pf != nullptr) {  // ask the L2 to fetch the NEXT kernel's weights: CTA share, split over the 16 warps
      const int64_t per_cta = (((p.pf_bytes + gridDim.x - 1) / gridDim.x) + 2047) & ~(int64_t)2047;
      const int64_t per_warp = per_cta / (kGroups * kSWarps);  // multiple of 128
      int64_t beg = (int64_t)blockIdx.x * per_cta + (int64_t)warp_all * per_warp;
      int64_t end = beg + per_warp;
      if (end > p.pf_bytes) end = p.pf_bytes & ~(int64_t)15;
      for (; beg < end; beg += 8192) prefetch_l2_bulk(p.pf + beg, (uint32_t)(end - beg < 8192 ? end - beg : 8192));
    }
  }
  for (int i = tid; i < 256 * 32; i += kSThreadsAll) {
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
    for (int c = tid; c < chunks; c += kSThreadsAll) {
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
      if (lane == 0) s_ssq[warp_all * 16 + b] = ss;
    }
  }
  if (p.norm_w) {
    __syncthreads();
    if (tid < B) {
      float tot = 0.f;
#pragma unroll
      for (int w = 0; w < kGroups * kSWarps; ++w) tot += s_ssq[w * 16 + tid];
      s_rstd[tid] = rsqrtf(tot / (float)p.K + p.eps);
    }
    __syncthreads();
    for (int b = 0; b < B; ++b) {
      const float rstd = s_rstd[b];
      for (int c = tid; c < chunks; c += kSThreadsAll) {  // same thread owns the same chunk as above
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
    for (int i = tid; i < B * p.nkb; i += kSThreadsAll) {
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

  // one k-block of both tiles: w0 / w1 = this lane's 16 B {row g bytes, row g+8 bytes}; s0 / s1 = scale words
  auto compute = [&](const uint4 w0, const uint4 w1, const uint2 s0, const uint2 s1, int kb) {
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
        const uint32_t wa = (j < 2) ? w0.x : w0.y, wb = (j < 2) ? w0.z : w0.w;
        const uint32_t a0 = LUT(wa, sel0), a2 = LUT(wa, sel1), a1 = LUT(wb, sel0), a3 = LUT(wb, sel1);
#pragma unroll
        for (int nb = 0; nb < NB; ++nb) mma16816(p0[nb], a0, a1, a2, a3, xb[nb][2 * j], xb[nb][2 * j + 1]);
      }
      {
        const uint32_t wa = (j < 2) ? w1.x : w1.y, wb = (j < 2) ? w1.z : w1.w;
        const uint32_t a0 = LUT(wa, sel0), a2 = LUT(wa, sel1), a1 = LUT(wb, sel0), a3 = LUT(wb, sel1);
#pragma unroll
        for (int nb = 0; nb < NB; ++nb) mma16816(p1[nb], a0, a1, a2, a3, xb[nb][2 * j], xb[nb][2 * j + 1]);
      }
    }
    if (INT4) {
      const __half2 h0a = *reinterpret_cast<const __half2*>(&s0.x), h0b = *reinterpret_cast<const __half2*>(&s0.y);
      const __half2 h1a = *reinterpret_cast<const __half2*>(&s1.x), h1b = *reinterpret_cast<const __half2*>(&s1.y);
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
      const float s0a = __uint_as_float(s0.x), s0b = __uint_as_float(s0.y);
      const float s1a = __uint_as_float(s1.x), s1b = __uint_as_float(s1.y);
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

  int stage = 0;
  uint32_t parity = 0;
  auto group_of = [&](int u) -> int {  // group whose range contains unit u
    int j = (int)(((int64_t)u * ngroups) / Wtot);
    while (j + 1 < ngroups && bnd(j + 1) <= u) ++j;
    while (j > 0 && bnd(j) > u) --j;
    return j;
  };

  SegCursor cs;
  cs.cur = u_begin; cs.end = u_end; cs.pair = 0; cs.a = cs.b = 0; cs.kb0 = cs.kb1 = 0;
  bool live = u_begin < u_end;
  if (live) seg_load(cs, p.upp, BPU, warp);
  for (; live; live = seg_next(cs, p.upp, BPU, warp)) {
    const int pair = cs.pair;
    const bool has1 = (p.swiglu ? pair : 2 * pair + 1) < p.ntiles;
    for (int cb = cs.kb0; cb < cs.kb1; cb += C) {
      const int nblk = min(C, cs.kb1 - cb);
      mbar_wait(my_bars + stage * 8, parity);
      const uint8_t* sp = my_ring + stage * STAGE;
#pragma unroll
      for (int b = 0; b < C; ++b) {
        if (b < nblk) {
          const int woff = (INT4 ? b * 512 : b * 576) + lane * 16;
          const int soff = (INT4 ? 1024 : b * 576 + 512) + g * 8;
          const uint4 w0 = *reinterpret_cast<const uint4*>(sp + woff);
          const uint2 s0 = *reinterpret_cast<const uint2*>(sp + soff);
          uint4 w1 = make_uint4(0u, 0u, 0u, 0u);
          uint2 s1 = make_uint2(0u, 0u);
          if (has1) {
            w1 = *reinterpret_cast<const uint4*>(sp + STAGE_TILE + woff);
            s1 = *reinterpret_cast<const uint2*>(sp + STAGE_TILE + soff);
          }
          compute(w0, w1, s0, s1, cb + b);
        }
      }
      __syncwarp();
      if (lane == 0) {  // the stage is drained: hand it back to the async proxy and refill it S chunks ahead
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        produce();
      }
      if (++stage == S) { stage = 0; parity ^= 1u; }
    }
    // ---- reduce the 8 k-slices of this group through shared memory ----------------------------------------------
    float* rbuf = red + grp * RED_GROUP + warp * (2 * 16 * COLS);
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
      *reinterpret_cast<float2*>(rbuf + g * COLS + 8 * nb + 2 * t) = make_float2(acc0[nb][0], acc0[nb][1]);
      *reinterpret_cast<float2*>(rbuf + (g + 8) * COLS + 8 * nb + 2 * t) = make_float2(acc0[nb][2], acc0[nb][3]);
      *reinterpret_cast<float2*>(rbuf + 16 * COLS + g * COLS + 8 * nb + 2 * t) = make_float2(acc1[nb][0], acc1[nb][1]);
      *reinterpret_cast<float2*>(rbuf + 16 * COLS + (g + 8) * COLS + 8 * nb + 2 * t) = make_float2(acc1[nb][2], acc1[nb][3]);
#pragma unroll
      for (int i = 0; i < 4; ++i) acc0[nb][i] = acc1[nb][i] = 0.f;
    }
    asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "n"(kSThreads) : "memory");
    const bool reducer = gtid < 16 * COLS;
    const int r = gtid / COLS, c = gtid % COLS;
    float v0 = 0.f, v1 = 0.f;
    if (reducer) {
      const float* rb = red + grp * RED_GROUP;
#pragma unroll
      for (int w = 0; w < kSWarps; ++w) {
        v0 += rb[w * (2 * 16 * COLS) + r * COLS + c];
        v1 += rb[w * (2 * 16 * COLS) + 16 * COLS + r * COLS + c];
      }
    }
    bool do_epi = true;
    if (cs.a != 0 || cs.b != p.upp) {  // the pair is cut by a group boundary (group-uniform): deterministic hand-off
      const int jfirst = group_of(pair * p.upp), jlast = group_of(pair * p.upp + p.upp - 1);
      const int slot = gidx - jfirst, ncontrib = jlast - jfirst + 1;
      float* part = p.ws_part + ((int64_t)pair * p.maxslots) * (2 * 16 * COLS);
      if (reducer) {
        __stcg(part + slot * (2 * 16 * COLS) + r * COLS + c, v0);
        __stcg(part + slot * (2 * 16 * COLS) + 16 * COLS + r * COLS + c, v1);
        __threadfence();
      }
      asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "n"(kSThreads) : "memory");
      if (gtid == 0) {
        const unsigned int old = atomicAdd(p.ws_cnt + pair, 1u);
        const bool last = old == (unsigned int)(ncontrib - 1);
        if (last) p.ws_cnt[pair] = 0u;  // self-cleaning for the next launch
        s_flag[grp] = last ? 1u : 0u;
      }
      asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "n"(kSThreads) : "memory");
      do_epi = s_flag[grp] != 0u;
      if (do_epi && reducer) {
        __threadfence();
        v0 = v1 = 0.f;
        for (int sl = 0; sl < ncontrib; ++sl) {  // fixed order: deterministic
          v0 += __ldcg(part + sl * (2 * 16 * COLS) + r * COLS + c);
          v1 += __ldcg(part + sl * (2 * 16 * COLS) + 16 * COLS + r * COLS + c);
        }
      }
    }
    if (do_epi && reducer && c < B) {
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
    asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "n"(kSThreads) : "memory");
  }
}

static std::atomic<int> g_stream_depth{0}, g_stream_occ{0}, g_stream_pf{1};

template <int NB, bool INT4>
static cudaError_t launch_stream(const StreamParams& p, int gx, size_t smem, cudaStream_t st) {
  static bool configured = false;  // per instantiation
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(qgemv_stream_kernel<NB, INT4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024));
    if (e != cudaSuccess) return e;
    configured = true;
  }
  return launch(qgemv_stream_kernel<NB, INT4>, dim3(gx), dim3(kSThreadsAll), smem, st, p);
}

}  // namespace b200

using namespace b200;

extern "C" {

size_t b200_qgemv_tiled_workspace_bytes(void) { return (size_t)32 << 20; }

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
  const size_t stage_bytes = 2 * (size_t)kChunkBlocks * (int4 ? 544 : 576);
  size_t fixed = kSLutBytes + 8 * (size_t)kGroups * kSWarps * kMaxStages +
                 sizeof(float) * ((size_t)kGroups * kSWarps * 2 * 16 * COLS + kGroups * kSWarps * 16 + 32);
  if (int4) fixed += sizeof(float) * (((size_t)a->batch * p.nkb + 3) & ~(size_t)3);
  fixed += ((size_t)a->batch * p.xstride * 2 + 127) & ~(size_t)127;
  fixed += 128;  // alignment slack of the dynamic shared-memory base
  int stages = g_stream_depth.load();
  if (stages < 2 || stages > kMaxStages) stages = 3;
  while (stages > 2 && fixed + (size_t)kGroups * kSWarps * stages * stage_bytes > 227 * 1024) --stages;
  const size_t smem = fixed + (size_t)kGroups * kSWarps * stages * stage_bytes;
  if (smem > 227 * 1024) {
    set_error("b200_qgemv_tiled: batch %d x K %lld needs %zu bytes of shared memory (> 227 KB)", a->batch, (long long)a->K, smem);
    return B200_ERR_UNSUPPORTED;
  }
  p.stages = stages;
  p.upp = int4 ? p.nkb / 2 : p.nkb;
  // stream-K: every group gets the same number of work units; keep >= 8 units per group so its 8 warps have work
  const int64_t Wtot = (int64_t)p.npairs * p.upp;
  int gx = sm_count();
  if ((int64_t)gx * kGroups * 8 > Wtot) gx = (int)((Wtot + kGroups * 8 - 1) / (kGroups * 8));
  const int ngroups = gx * kGroups;
  p.maxslots = (ngroups + p.npairs - 1) / p.npairs + 1;
  const size_t cnt_bytes = ((size_t)p.npairs * sizeof(unsigned int) + 255) & ~(size_t)255;
  const size_t need = cnt_bytes + (size_t)p.npairs * p.maxslots * 2 * 16 * COLS * sizeof(float);
  B200_CHECK_ARG(a->workspace && a->workspace_bytes >= need, "b200_qgemv_tiled: workspace too small (%zu < %zu)", (size_t)a->workspace_bytes, need);
  B200_CHECK_ARG((reinterpret_cast<uintptr_t>(a->workspace) & 255) == 0, "b200_qgemv_tiled: workspace must be 256-byte aligned");
  p.ws_cnt = reinterpret_cast<unsigned int*>(a->workspace);
  p.ws_part = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(a->workspace) + cnt_bytes);
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e;
  if (int4) e = NB == 2 ? launch_stream<2, true>(p, gx, smem, st) : launch_stream<1, true>(p, gx, smem, st);
  else e = NB == 2 ? launch_stream<2, false>(p, gx, smem, st) : launch_stream<1, false>(p, gx, smem, st);
  if (e != cudaSuccess) {
    set_error("b200_qgemv_tiled: launch failed: %s", cudaGetErrorString(e));
    return B200_ERR_CUDA;
  }
  return B200_OK;
}

}  // extern "C"
