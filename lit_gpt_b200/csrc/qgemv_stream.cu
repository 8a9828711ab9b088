// Streaming fused dequant-GEMV over the TILED weight layout (csrc/tile.cu), batch 1..16.
//
// Replaces bnb.nn.Linear4bit.forward at lit_gpt/model.py:619,656,713,714,716,519.
//
// One persistent CTA per SM owns a contiguous range of 16-row tiles (whole output rows: no split-K across CTAs, no
// workspace, no atomics).  It walks them two tiles at a time (rows u..u+31, or, for SwiGLU, tile p of the gate and
// tile p of the up matrix); the 16 consumer warps split K between them.  A 17th PRODUCER warp streams every
// consumer warp's weights through that warp's private ring of shared-memory stages with cp.async.bulk (TMA) copies
// signalled on mbarriers (full / empty), so bytes in flight are bounded by shared memory, not by the register
// scoreboard, and run ahead across tile boundaries and, under programmatic dependent launch, across the previous
// kernel's tail.  The producer also issues cp.async.bulk.prefetch.L2 for a caller-given range (the NEXT kernel's
// weights) so HBM stays busy across the kernel boundary.
// Dequant: ONE PRMT turns a packed byte into the complete shared-memory address of a per-lane replicated
// 256-entry byte -> bf16x2 table (the table sits at a 64 KB-aligned shared address so no add is needed), one
// conflict-free LDS returns two code values in mma.sync A-fragment order; mma.sync m16n8k16 (bf16 x bf16 -> fp32)
// does the MACs against activation B-fragments read from shared memory (shared by the two tiles); the per-64
// absmax (or int4 {scale, zero}) is applied in fp32 per k-block.
// Prologue: residual add + RMSNorm of the activation (lit_gpt/rmsnorm.py:19-25, model.py:591) into shared memory.
// Epilogue: bias / residual add / SwiGLU with the reference's bf16 rounding points (lit_gpt/model.py:712-716).
#include "common.cuh"
#include "qgemv_v4.cuh"

namespace b200 {

struct StreamParams {
  const uint8_t* tA;   // tiled matrix (tile.cu layout)
  const uint8_t* tB;   // second tiled matrix for SwiGLU (else == tA)
  int64_t N, K;
  int nkb;             // K / 64
  int ntiles;          // ceil(N / 16)
  int nunits;          // work units: SwiGLU: ntiles (gate tile + up tile each), else ntiles
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
  int red_bufs;        // reduction buffers (host-chosen: 2 when the ring still gets >= 3 stages, else 1)
  int debug;           // 1: consumers skip the dequant/MMA work (streaming ceiling measurement only)
  const uint8_t* pf;   // L2 prefetch range (next kernel's weights), may be null
  int64_t pf_bytes;
  // tensor-parallel hand-off over peer memory (b200_tp_handoff): 32-bit packets tag << 16 | bf16
  int tp_world;
  uint32_t* tp_peer_y[8];        // producer: this rank's slot on every rank
  const uint32_t* tp_send_seq;   // producer: local sequence counter of the buffer written
  uint32_t* err;             // error word (workspace[0]): 2 = a tensor-parallel peer did not deliver within ~2 s
  const uint32_t* tp_slots;      // consumer: local slots (rank r at + r * tp_slot_stride packets)
  int64_t tp_slot_stride;
  uint32_t* tp_recv_seq;         // consumer: local sequence counter of the buffer read
  uint32_t* tp_recv_done;        // consumer: local CTA counter
};

constexpr int kSThreads = 256;
constexpr int kSWarps = 8;
constexpr int kSLutBytes = 256 * 256;  // entry e, lane l at byte e*256 + l*4

__device__ __forceinline__ void mma16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm(  // not volatile: a pure function of its operands, ptxas may interleave it with the table look-ups
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

__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

constexpr int kCW = 16;                          // consumer warps (they split K)
constexpr int kSThreadsAll = (kCW + 1) * 32;     // + 1 producer warp
constexpr int kConsumers = kCW * 32;
constexpr int kMaxStages = 6;
constexpr int kDynSmem = 227 * 1024;

// Shared memory is addressed by ABSOLUTE shared-window addresses so that the dequant table can sit at a 64 KB
// aligned address L (then `byte << 8 | lane << 2 | L` is assembled by one PRMT):
//   [b0 .. )        mbarriers, reduction buffer, norm scratch, int4 activation sums, activation rows x_s
//   [.. L)          ring stages (region A)
//   [L, L + 64 KB)  table
//   [L + 64 KB ..)  ring stages (region C)
template <int NB, bool INT4>
__global__ void __launch_bounds__(kSThreadsAll, 1) qgemv_stream_kernel(const StreamParams p) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  constexpr int COLS = 8 * NB;
  constexpr int REC = INT4 ? 1088 : 576;         // one k-unit (NF4: 64 k, INT4: 128 k) of one tile
  constexpr int KPU = INT4 ? 2 : 1;              // k-blocks per unit
  constexpr int STAGE_TILE = kCW * REC;          // 16 consecutive units of one tile
  constexpr int STAGE = 2 * STAGE_TILE;          // two tiles
  const uint32_t b0 = smem_u32(smem_raw);
  uint8_t* const sb = smem_raw - b0;             // generic pointer to shared-window address 0

  uint32_t off = (b0 + 127u) & ~127u;
  const uint32_t bars_a = off;          off += 2 * kMaxStages * 8;                      // full[6], empty[6]
  off = (off + 127u) & ~127u;
  const int RED_BUFS = p.red_bufs;            // 2: double-buffered (one barrier per step); 1 when shared memory is short
  const uint32_t red_a = off;           off += RED_BUFS * kCW * 2 * 16 * COLS * 4;      // [buffers][16 warps][2 tiles][16][COLS]
  const uint32_t ssq_a = off;           off += kCW * 16 * 4;                            // [16 warps][16]
  const uint32_t rstd_a = off;          off += 16 * 4;
  const uint32_t xsum_a = off;          off += INT4 ? (((uint32_t)(p.batch * p.nkb) + 3u) & ~3u) * 4 : 0;
  const uint32_t xs_a = off;            off += (((uint32_t)p.batch * p.xstride * 2) + 127u) & ~127u;
  const uint32_t ringA_a = off;
  // int4 dequantises arithmetically: no table, the whole remainder is ring
  const uint32_t lut_a = INT4 ? off : (off + 0xFFFFu) & ~0xFFFFu;                       // 64 KB aligned
  const uint32_t ringC_a = INT4 ? off : lut_a + 0x10000u;
  const uint32_t end_a = b0 + kDynSmem;
  const int nA = (int)((lut_a - ringA_a) / STAGE);
  const int nC = ringC_a < end_a ? (int)((end_a - ringC_a) / STAGE) : 0;
  const int S = min(kMaxStages, nA + nC);
  if (S < 2 || ringC_a > end_a) __trap();        // the host sized the launch so that this cannot happen
  auto stage_addr = [&](int s) -> uint32_t { return s < nC ? ringC_a + s * STAGE : ringA_a + (s - nC) * STAGE; };

  float* red = reinterpret_cast<float*>(sb + red_a);
  float* s_ssq = reinterpret_cast<float*>(sb + ssq_a);
  float* s_rstd = reinterpret_cast<float*>(sb + rstd_a);
  float* xsum_s = reinterpret_cast<float*>(sb + xsum_a);
  __nv_bfloat16* x_s = reinterpret_cast<__nv_bfloat16*>(sb + xs_a);
  uint32_t* lut = reinterpret_cast<uint32_t*>(sb + lut_a);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int B = p.batch;
  const int nku = p.nkb / KPU;                                  // k-units per tile
  const int nrounds = (nku + kCW - 1) / kCW;                    // ring stages consumed per step
  // this CTA's contiguous range of work units; a step takes two tiles (non-SwiGLU) or one gate/up pair (SwiGLU)
  const int u0 = (int)(((int64_t)p.nunits * blockIdx.x) / gridDim.x);
  const int u1 = (int)(((int64_t)p.nunits * (blockIdx.x + 1)) / gridDim.x);
  const int ustep = p.swiglu ? 1 : 2;

  if (tid == 0) {
    for (int s = 0; s < kMaxStages; ++s) { mbar_init(bars_a + s * 8, 1); mbar_init(bars_a + (kMaxStages + s) * 8, kCW); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  if (warp == kCW) {
    // ================================ PRODUCER WARP (one elected lane) ==========================================
    if (lane == 0) {
      const uint64_t pol = make_evict_first_policy();
      int n = 0, ps = 0;          // global stage counter, ring slot
      uint32_t pph = 1;           // parity to wait for on the slot's EMPTY barrier (first lap: no wait)
      auto prefetch_next = [&]() {  // ask the L2 to fetch the NEXT kernel's weights (this CTA's share)
        if (p.pf == nullptr) return;
        const int64_t per_cta = (((p.pf_bytes + gridDim.x - 1) / gridDim.x) + 127) & ~(int64_t)127;
        int64_t beg = (int64_t)blockIdx.x * per_cta, end = beg + per_cta;
        if (end > p.pf_bytes) end = p.pf_bytes & ~(int64_t)15;
        for (; beg < end; beg += 16384) prefetch_l2_bulk(p.pf + beg, (uint32_t)(end - beg < 16384 ? end - beg : 16384));
      };
      for (int u = u0; u < u1; u += ustep) {
        const int tile0 = u, tile1 = p.swiglu ? u : u + 1;
        const bool has1 = p.swiglu || tile1 < u1;
        for (int r = 0; r < nrounds; ++r, ++n) {
          if (n >= S) mbar_wait(bars_a + (kMaxStages + ps) * 8, pph);
          const int nu = min(kCW, nku - r * kCW);
          const uint32_t bytes = (uint32_t)nu * REC;
          const uint32_t bar = bars_a + ps * 8, dst = stage_addr(ps);
          mbar_expect_tx(bar, has1 ? 2 * bytes : bytes);
          bulk_g2s(dst, p.tA + ((int64_t)tile0 * nku + r * kCW) * REC, bytes, bar, pol);
          if (has1) bulk_g2s(dst + STAGE_TILE, p.tB + ((int64_t)tile1 * nku + r * kCW) * REC, bytes, bar, pol);
          if (n + 1 == S) prefetch_next();  // once this kernel's own ring is full
          if (++ps == S) { ps = 0; pph ^= 1u; }
        }
      }
      if (n < S) prefetch_next();
    }
    griddep_launch();
    return;
  }
  // ================================== CONSUMER WARPS ===============================================================
  if (!INT4) {
    // table entry e = warp + 16 k, lane slot `lane`: element 0 (low half) = code[e >> 4] = code[k], element 1 = code[e & 15] =
    // code[warp] (even element = HIGH nibble).  Sixteen independent constant loads + stores per thread, fully unrolled: the
    // fill is on every SM's critical path between two kernels (a rolled loop with two dependent indexed constant loads per
    // iteration cost ~0.8 us per launch)
    const float* cb = p.qtype == B200_Q_NF4 ? kNF4 : kFP4;
    const float mine = cb[warp];
#pragma unroll
    for (int k = 0; k < 16; ++k) lut[(warp + 16 * k) * 64 + lane] = pack_bf16(cb[k], mine);
  }
  // RMSNorm weights are constants: fetch this thread's first chunks BEFORE waiting for the previous kernel, so their
  // global-load latency is off the critical path between the two kernels
  const int chunks = (int)(p.K >> 3);
  uint4 nwreg[3] = {make_uint4(0u, 0u, 0u, 0u), make_uint4(0u, 0u, 0u, 0u), make_uint4(0u, 0u, 0u, 0u)};
  if (p.norm_w) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const int c = tid + i * kConsumers;
      if (c < chunks) nwreg[i] = *reinterpret_cast<const uint4*>(p.norm_w + c * 8);
    }
  }
  griddep_launch();
  griddep_wait();

  // ---- tensor-parallel consumer: the branch output is the sum of every rank's partial, written into this rank's
  // slots by the peers' epilogues (and its own) as self-validating packets; the expected tag comes from the buffer's
  // local sequence counter (last bumped at least two launches ago)
  uint32_t rtag = 0;
  if (p.tp_slots) rtag = (__ldcg(p.tp_recv_seq) % 65535u) + 1u;
  // ---- activation prologue: (x + delta) -> optional RMSNorm -> bf16 rows in shared memory ---------------------
  const bool designated = blockIdx.x == 0;
  for (int b = 0; b < B; ++b) {
    float ss = 0.f;
    for (int c = tid; c < chunks; c += kConsumers) {
      uint4 v = *reinterpret_cast<const uint4*>(p.x + (int64_t)b * p.K + c * 8);
      if (p.delta || p.tp_slots) {
        uint4 dv;
        if (p.tp_slots) {  // sum of the ranks' partials in rank order (fp32), rounded once to bf16 like an all-reduced tensor
          float acc8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
          for (int r = 0; r < p.tp_world; ++r) {
            const uint32_t* sp = p.tp_slots + r * p.tp_slot_stride + (int64_t)b * p.K + c * 8;
            uint4 pa, pb;
            const long long t0 = clock64();
            for (;;) {   // poll until all eight packets carry this hand-off's tag
              asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(pa.x), "=r"(pa.y), "=r"(pa.z), "=r"(pa.w) : "l"(sp) : "memory");
              asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(pb.x), "=r"(pb.y), "=r"(pb.z), "=r"(pb.w) : "l"(sp + 4) : "memory");
              const bool ok = (pa.x >> 16) == rtag && (pa.y >> 16) == rtag && (pa.z >> 16) == rtag && (pa.w >> 16) == rtag &&
                              (pb.x >> 16) == rtag && (pb.y >> 16) == rtag && (pb.z >> 16) == rtag && (pb.w >> 16) == rtag;
              if (ok) break;
              if (clock64() - t0 > 4000000000LL) {  // a peer died: raise the error word (host: b200_qgemv_tiled_error) and go on; no hang, no trap
                if (p.err) *reinterpret_cast<volatile uint32_t*>(p.err) = 2u;
                break;
              }
            }
            acc8[0] += bf16lo(pa.x); acc8[1] += bf16lo(pa.y); acc8[2] += bf16lo(pa.z); acc8[3] += bf16lo(pa.w);
            acc8[4] += bf16lo(pb.x); acc8[5] += bf16lo(pb.y); acc8[6] += bf16lo(pb.z); acc8[7] += bf16lo(pb.w);
          }
          dv = make_uint4(pack_bf16(acc8[0], acc8[1]), pack_bf16(acc8[2], acc8[3]), pack_bf16(acc8[4], acc8[5]), pack_bf16(acc8[6], acc8[7]));
        } else {
          dv = *reinterpret_cast<const uint4*>(p.delta + (int64_t)b * p.K + c * 8);
        }
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
    asm volatile("bar.sync 1, %0;" ::"n"(kConsumers) : "memory");
    if (tid < B) {
      float tot = 0.f;
#pragma unroll
      for (int w = 0; w < kCW; ++w) tot += s_ssq[w * 16 + tid];
      s_rstd[tid] = rsqrtf(tot / (float)p.K + p.eps);
    }
    asm volatile("bar.sync 1, %0;" ::"n"(kConsumers) : "memory");
    for (int b = 0; b < B; ++b) {
      const float rstd = s_rstd[b];
      int ci = 0;
      for (int c = tid; c < chunks; c += kConsumers, ++ci) {  // same thread owns the same chunk as above
        uint4 v = *reinterpret_cast<uint4*>(x_s + b * p.xstride + c * 8);
        const uint4 nw = ci == 0 ? nwreg[0] : ci == 1 ? nwreg[1] : ci == 2 ? nwreg[2] : *reinterpret_cast<const uint4*>(p.norm_w + c * 8);
        // weight * (x * rstd), fp32, one cast (rmsnorm.py:24-25)
        v.x = pack_bf16(bf16lo(nw.x) * (bf16lo(v.x) * rstd), bf16hi(nw.x) * (bf16hi(v.x) * rstd));
        v.y = pack_bf16(bf16lo(nw.y) * (bf16lo(v.y) * rstd), bf16hi(nw.y) * (bf16hi(v.y) * rstd));
        v.z = pack_bf16(bf16lo(nw.z) * (bf16lo(v.z) * rstd), bf16hi(nw.z) * (bf16hi(v.z) * rstd));
        v.w = pack_bf16(bf16lo(nw.w) * (bf16lo(v.w) * rstd), bf16hi(nw.w) * (bf16hi(v.w) * rstd));
        *reinterpret_cast<uint4*>(x_s + b * p.xstride + c * 8) = v;
      }
    }
  }
  asm volatile("bar.sync 1, %0;" ::"n"(kConsumers) : "memory");  // x_s and the LUT are published
  if (p.tp_slots && tid == 0) {  // last CTA to have read the slots bumps the buffer's sequence counter for its next use
    const uint32_t old = atomicAdd(p.tp_recv_done, 1u);
    if (old == gridDim.x - 1) {
      *p.tp_recv_done = 0u;
      *p.tp_recv_seq = __ldcg(p.tp_recv_seq) + 1u;
    }
  }
  if (INT4) {       // per (batch row, k-block) activation sums for the zero-point term
    xsum_blocks<kConsumers>(x_s, p.xstride, B, p.nkb, xsum_s, tid);
    asm volatile("bar.sync 1, %0;" ::"n"(kConsumers) : "memory");
  }

  // table address = L | byte << 8 | lane << 2, assembled by ONE PRMT (L is 64 KB aligned): picks lane_full.b0,
  // word.b_i, lane_full.b2, lane_full.b3
  const uint32_t lane_full = lut_a | ((uint32_t)lane << 2);
  auto LUT = [&](uint32_t word, uint32_t sel) -> uint32_t {
    uint32_t v;
    asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(__byte_perm(word, lane_full, sel)));  // table is read-only: not volatile
    return v;
  };

  // int4 unpack constants kept in registers (opaque to constant propagation so that they stay register operands)
  uint32_t i4mask = 0x000F000Fu, i4magic = 0x43004300u;
  asm volatile("" : "+r"(i4mask), "+r"(i4magic));
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
      if constexpr (INT4) {
        // no table: the tiled words hold even codes in the low and odd codes in the high half-word, so one shift +
        // one LOP3 yields the bf16 pair (128 + e_2p, 128 + e_2p+1); the +128 joins the zero-point term below
        const int sh = 8 * (j & 1);
        const uint32_t ua = (j < 2) ? w0.x : w0.y, ub = (j < 2) ? w0.z : w0.w, va = (j < 2) ? w1.x : w1.y, vb = (j < 2) ? w1.z : w1.w;
        // (w >> s) & 0x000F000F | 0x43004300 as ONE three-input LOP3 (mask and magic in registers: with two immediates
        // the compiler emits two LOP3 per word)
        auto ex = [&](uint32_t w, int s) -> uint32_t {
          uint32_t r;
          asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(r) : "r"(s ? w >> s : w), "r"(i4mask), "r"(i4magic));
          return r;
        };
        const uint32_t a0 = ex(ua, sh), a1 = ex(ub, sh), a2 = ex(ua, sh + 4), a3 = ex(ub, sh + 4);
        const uint32_t b0 = ex(va, sh), b1 = ex(vb, sh), b2 = ex(va, sh + 4), b3 = ex(vb, sh + 4);
#pragma unroll
        for (int nb = 0; nb < NB; ++nb) {
          mma16816(p0[nb], a0, a1, a2, a3, xb[nb][2 * j], xb[nb][2 * j + 1]);
          mma16816(p1[nb], b0, b1, b2, b3, xb[nb][2 * j], xb[nb][2 * j + 1]);
        }
        continue;
      }
      const uint32_t sel0 = (j & 1) ? 0x7624u : 0x7604u, sel1 = (j & 1) ? 0x7634u : 0x7614u;
      {
        const uint32_t wa = (j < 2) ? w0.x : w0.y, wb = (j < 2) ? w0.z : w0.w;
        const uint32_t a0 = LUT(wa, sel0), a2 = LUT(wa, sel1), a1 = LUT(wb, sel0), a3 = LUT(wb, sel1);
#pragma unroll
        for (int nb = 0; nb < NB; ++nb) mma16816(p0[nb], a0, a1, a2, a3, xb[nb][2 * j], xb[nb][2 * j + 1]);
      }
      {  // tile 1 is computed even when absent (stale stage bytes, result never stored): no branch in the hot block
        const uint32_t wa = (j < 2) ? w1.x : w1.y, wb = (j < 2) ? w1.z : w1.w;
        const uint32_t a0 = LUT(wa, sel0), a2 = LUT(wa, sel1), a1 = LUT(wb, sel0), a3 = LUT(wb, sel1);
#pragma unroll
        for (int nb = 0; nb < NB; ++nb) mma16816(p1[nb], a0, a1, a2, a3, xb[nb][2 * j], xb[nb][2 * j + 1]);
      }
    }
    if (INT4) {
      const __half2 h0a = *reinterpret_cast<const __half2*>(&s0.x), h0b = *reinterpret_cast<const __half2*>(&s0.y);
      const __half2 h1a = *reinterpret_cast<const __half2*>(&s1.x), h1b = *reinterpret_cast<const __half2*>(&s1.y);
      const float s0a = __low2float(h0a), z0a = __high2float(h0a) + 128.f, s0b = __low2float(h0b), z0b = __high2float(h0b) + 128.f;
      const float s1a = __low2float(h1a), z1a = __high2float(h1a) + 128.f, s1b = __low2float(h1b), z1b = __high2float(h1b) + 128.f;
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

  const uint32_t stag = p.tp_send_seq ? (__ldcg(p.tp_send_seq) % 65535u) + 1u : 0u;  // after griddep_wait: last bumped >= 2 launches ago
  auto store_out = [&](int64_t n, int c, float v) {  // STORE / RESIDUAL epilogues for one (row, batch column)
    float h = v + (p.bias ? __bfloat162float(p.bias[n]) : 0.f);
    if (p.epilogue == B200_EPI_RESIDUAL) h = round_bf16(h) + __bfloat162float(p.residual[(int64_t)c * p.N + n]);
    const __nv_bfloat16 hb = __float2bfloat16_rn(h);
    if (p.y) p.y[(int64_t)c * p.N + n] = hb;
    if (p.tp_send_seq) {  // row-parallel partial: one self-validating packet straight into every rank's slot over NVLink
      const uint32_t pk = (stag << 16) | (uint32_t)bf16_bits(h);
      for (int r = 0; r < p.tp_world; ++r)
        asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p.tp_peer_y[r] + (int64_t)c * p.N + n), "r"(pk) : "memory");
    }
  };


  // Register-level software pipeline over the flattened (step, round) sequence: the shared-memory stage of round
  // i+1 is read into registers before round i is computed, and a stage is handed back to the producer one round
  // after it was read (its loads have certainly completed by then).
  constexpr int RED_ONE = kCW * 2 * 16 * COLS;
  const int nsteps = (u1 - u0 + ustep - 1) / ustep;
  const int total = nsteps * nrounds;
  uint4 c0[KPU], c1[KPU], n0[KPU], n1[KPU];
  uint2 cs0, cs1, ns0, ns1;
  int ls = 0;              // ring slot of the next stage to read
  uint32_t lph = 0;        // its FULL-barrier parity
  auto read_stage = [&](int r, uint4 (&a0)[KPU], uint4 (&a1)[KPU], uint2& b0v, uint2& b1v) {
    mbar_wait(bars_a + ls * 8, lph);
    const int ku = r * kCW + warp;
    if (ku < nku) {
      const uint32_t sp = stage_addr(ls) + warp * REC;
      const uint32_t sa = sp + KPU * 512 + g * 8;
#pragma unroll
      for (int b = 0; b < KPU; ++b) {
        const uint32_t wa = sp + b * 512 + lane * 16;
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(a0[b].x), "=r"(a0[b].y), "=r"(a0[b].z), "=r"(a0[b].w) : "r"(wa));
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(a1[b].x), "=r"(a1[b].y), "=r"(a1[b].z), "=r"(a1[b].w) : "r"(wa + STAGE_TILE));
      }
      asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(b0v.x), "=r"(b0v.y) : "r"(sa));
      asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(b1v.x), "=r"(b1v.y) : "r"(sa + STAGE_TILE));
    }
    if (++ls == S) { ls = 0; lph ^= 1u; }
  };
#pragma unroll
  for (int b = 0; b < KPU; ++b) c0[b] = c1[b] = n0[b] = n1[b] = make_uint4(0u, 0u, 0u, 0u);
  cs0 = cs1 = ns0 = ns1 = make_uint2(0u, 0u);

  const int nku_c = p.debug == 1 ? 0 : nku;  // debug: stream only, no dequant/MMA work
  int rs = 0;  // ring slot to release next (trails the read slot by one round)
  int u = u0, r = 0, it = 0;
  bool has1 = p.swiglu || u + 1 < u1;
  if (total > 0) read_stage(0, c0, c1, cs0, cs1);
  // One round; the stage registers ping-pong between two sets (the loop below is unrolled by two) so that handing the
  // prefetched stage over costs no register moves.
  auto do_round = [&](const int i, uint4 (&pa0)[KPU], uint4 (&pa1)[KPU], uint2& pas0, uint2& pas1, uint4 (&pb0)[KPU], uint4 (&pb1)[KPU],
                      uint2& pbs0, uint2& pbs1) {
    int nu_ = u, nr = r + 1;  // position of round i+1
    if (nr == nrounds) { nr = 0; nu_ = u + ustep; }
    const bool nhas1 = p.swiglu || nu_ + 1 < u1;
    if (i + 1 < total) read_stage(nr, pb0, pb1, pbs0, pbs1);
    const int ku = r * kCW + warp;
    if (ku < nku_c) {
#pragma unroll
      for (int b = 0; b < KPU; ++b) compute(pa0[b], pa1[b], pas0, pas1, ku * KPU + b);
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(bars_a + (kMaxStages + rs) * 8);  // round i's stage: read one iteration ago
    if (++rs == S) rs = 0;
    if (nr == 0) {
      // ---- end of a step: reduce the 16 k-slices through shared memory (double-buffered: one barrier per step) ---
      float* rbuf = red + (RED_BUFS == 2 ? (it & 1) : 0) * RED_ONE + warp * (2 * 16 * COLS);
#pragma unroll
      for (int nb = 0; nb < NB; ++nb) {
        *reinterpret_cast<float2*>(rbuf + g * COLS + 8 * nb + 2 * t) = make_float2(acc0[nb][0], acc0[nb][1]);
        *reinterpret_cast<float2*>(rbuf + (g + 8) * COLS + 8 * nb + 2 * t) = make_float2(acc0[nb][2], acc0[nb][3]);
        *reinterpret_cast<float2*>(rbuf + 16 * COLS + g * COLS + 8 * nb + 2 * t) = make_float2(acc1[nb][0], acc1[nb][1]);
        *reinterpret_cast<float2*>(rbuf + 16 * COLS + (g + 8) * COLS + 8 * nb + 2 * t) = make_float2(acc1[nb][2], acc1[nb][3]);
#pragma unroll
        for (int k = 0; k < 4; ++k) acc0[nb][k] = acc1[nb][k] = 0.f;
      }
      asm volatile("bar.sync 1, %0;" ::"n"(kConsumers) : "memory");
      if (tid < 2 * 16 * COLS) {  // one thread per (tile, row, batch column)
        float* rb = red + (RED_BUFS == 2 ? (it & 1) : 0) * RED_ONE;
        const int tl = tid / (16 * COLS), rr = (tid / COLS) % 16, c = tid % COLS;
        float v = 0.f;
#pragma unroll
        for (int w = 0; w < kCW; ++w) v += rb[w * (2 * 16 * COLS) + tl * 16 * COLS + rr * COLS + c];
        if (p.swiglu) {
          // gate (tile 0) and up (tile 1) values of the same row meet in the gate thread through shared memory
          if (tl == 1) rb[16 * COLS + rr * COLS + c] = v;  // warp 0's tile-1 slot: only this thread ever reads it
          asm volatile("bar.sync 2, %0;" ::"n"(2 * 16 * COLS) : "memory");
          if (tl == 0 && c < B) {
            const int64_t nrow = (int64_t)u * 16 + rr;
            if (nrow < p.N) {
              const float a = round_bf16(v + (p.bias ? __bfloat162float(p.bias[nrow]) : 0.f));
              const float b = round_bf16(rb[16 * COLS + rr * COLS + c] + (p.bias2 ? __bfloat162float(p.bias2[nrow]) : 0.f));
              p.y[(int64_t)c * p.N + nrow] = __float2bfloat16_rn(silu_bf16_s(a) * b);
            }
          }
        } else if (c < B && (tl == 0 || has1)) {
          const int64_t nrow = (int64_t)(u + tl) * 16 + rr;
          if (nrow < p.N) store_out(nrow, c, v);
        }
      }
      if (RED_BUFS == 1) asm volatile("bar.sync 1, %0;" ::"n"(kConsumers) : "memory");
      ++it;
    }
    u = nu_; r = nr; has1 = nhas1;
  };
  for (int i = 0; i < total; i += 2) {
    do_round(i, c0, c1, cs0, cs1, n0, n1, ns0, ns1);
    if (i + 1 < total) do_round(i + 1, n0, n1, ns0, ns1, c0, c1, cs0, cs1);
  }
}

static std::atomic<int> g_nb_min_batch{2};   // plain rows from this batch on run the tile-per-warp kernel (qgemv_nb.cu); 17 = never
static std::atomic<int> g_stream_depth{0}, g_stream_occ{0}, g_stream_pf{0};  // L2 prefetch of the next launch's weights: off by default (no measured gain, extra DRAM traffic when lines are evicted before use)

template <int NB, bool INT4>
static cudaError_t launch_stream(const StreamParams& p, int gx, cudaStream_t st) {
  static PerDeviceOnce once;  // per instantiation
  if (once.need()) {
    cudaError_t e = cudaFuncSetAttribute(qgemv_stream_kernel<NB, INT4>, cudaFuncAttributeMaxDynamicSharedMemorySize, kDynSmem);
    if (e != cudaSuccess) return e;
    once.done();
  }
  return launch(qgemv_stream_kernel<NB, INT4>, dim3(gx), dim3(kSThreadsAll), (size_t)kDynSmem, st, p);
}

}  // namespace b200

using namespace b200;

extern "C" {

size_t b200_qgemv_tiled_workspace_bytes(void) { return qgemv_v4_workspace_bytes() + qgemv_nb_workspace_bytes(); }  // stream-K partial sums + flags (self-cleaning)

int b200_qgemv_tiled_tune(int depth, int ctas_per_sm, int l2_prefetch) {
  g_stream_depth = depth;
  g_stream_occ = ctas_per_sm;
  // development knobs: 10 = never split entities, 11 = always, 12 = no register-resident activations
  qgemv_v4_tune(ctas_per_sm == 11 ? 1 : ctas_per_sm == 10 ? 0 : -1, ctas_per_sm == 12 ? 0 : 1);
  qgemv_v4_tune_occ2(ctas_per_sm == 13 ? 1 : 0);
  qgemv_nb_tune_ru(ctas_per_sm >= 41 && ctas_per_sm <= 44 ? ctas_per_sm - 40 : 4);   // 41 / 42: at most 1 / 2 records per step and tile
  g_nb_min_batch = (ctas_per_sm >= 22 && ctas_per_sm <= 37) ? ctas_per_sm - 20 : 2;   // 22..36: batched kernel from 2..16 rows on, 37: off
  qgemv_v4_tune_kvhint(ctas_per_sm == 14 ? 0 : 1);   // 14: ignore the K / V cache L2 hint   // 13: int4 batch-1 as two co-resident CTAs per SM (experiment)
  g_stream_pf = l2_prefetch;
  return B200_OK;
}

int b200_qgemv_tiled_trace(void* buf, int slots) {
  qgemv_v4_set_trace(buf, slots);
  return B200_OK;
}

// dry = true: validate the shape against the kernels' plans only (no pointer checks, no launch)
static int qgemv_tiled_impl(const b200_qgemv_tiled_args* a, b200_stream_t stream, bool dry) {
  B200_CHECK_ARG(a != nullptr, "b200_qgemv_tiled: null args");
  B200_CHECK_ARG(dry || (a->tiled && a->x && (a->y || (a->tp && a->tp->send_seq))), "b200_qgemv_tiled: null tiled/x/y pointer");
  B200_CHECK_ARG(a->N > 0 && a->K > 0, "b200_qgemv_tiled: N and K must be positive (N=%lld K=%lld)", (long long)a->N, (long long)a->K);
  B200_CHECK_ARG(a->batch >= 1 && a->batch <= 16, "b200_qgemv_tiled: batch must be in [1,16] (got %d); use b200_qgemm for prefill", a->batch);
  B200_CHECK_ARG(a->qtype == B200_Q_NF4 || a->qtype == B200_Q_FP4 || a->qtype == B200_Q_INT4, "b200_qgemv_tiled: bad qtype %d", a->qtype);
  B200_CHECK_ARG(a->epilogue >= B200_EPI_STORE && a->epilogue <= B200_EPI_SWIGLU, "b200_qgemv_tiled: bad epilogue %d", a->epilogue);
  B200_CHECK_ARG(a->epilogue != B200_EPI_SWIGLU || a->tiled2, "b200_qgemv_tiled: SWIGLU needs tiled2");
  B200_CHECK_ARG(a->epilogue != B200_EPI_RESIDUAL || a->residual, "b200_qgemv_tiled: RESIDUAL epilogue needs residual");
  B200_CHECK_ARG(a->x_out == nullptr || a->x_out != a->x, "b200_qgemv_tiled: x_out must not alias x");
  const bool int4 = a->qtype == B200_Q_INT4;
  const int fold = a->fold > 1 ? a->fold : 1;
  if (fold > 1) {
    B200_CHECK_ARG((fold == 2 || fold == 4) && !int4 && a->N % fold == 0 && a->K % 8 == 0 && a->batch * fold <= 16,
                   "b200_qgemv_tiled: fold %d needs nf4 / fp4, N %% fold == 0, K %% 8 == 0 and batch * fold <= 16", fold);
    B200_CHECK_ARG(a->epilogue != B200_EPI_SWIGLU && !a->norm_weight && !a->delta && !a->x_out && !(a->tp && a->tp->slots),
                   "b200_qgemv_tiled: a folded matrix supports the STORE / RESIDUAL epilogues only (no norm, delta, x_out, tp consumer)");
  }
  B200_CHECK_ARG((a->K * fold) % (int4 ? 128 : 64) == 0, "b200_qgemv_tiled: K%s must be a multiple of %d (got %lld)", fold > 1 ? " * fold" : "",
                 int4 ? 128 : 64, (long long)a->K);
  B200_CHECK_ARG(((reinterpret_cast<uintptr_t>(a->x) | reinterpret_cast<uintptr_t>(a->tiled) | reinterpret_cast<uintptr_t>(a->tiled2) |
                   reinterpret_cast<uintptr_t>(a->delta) | reinterpret_cast<uintptr_t>(a->x_out) |
                   reinterpret_cast<uintptr_t>(a->norm_weight) | reinterpret_cast<uintptr_t>(a->prefetch)) & 15) == 0,
                 "b200_qgemv_tiled: pointers must be 16-byte aligned");

  if (a->tp) {
    const b200_tp_handoff* h = a->tp;
    B200_CHECK_ARG(h->world >= 2 && h->world <= 8 && h->rank >= 0 && h->rank < h->world, "b200_qgemv_tiled: tp world %d / rank %d out of range", h->world, h->rank);
    B200_CHECK_ARG(!(h->slots && a->delta), "b200_qgemv_tiled: tp slots replace delta (pass delta = NULL)");
    B200_CHECK_ARG(!(h->send_seq && a->epilogue == B200_EPI_SWIGLU), "b200_qgemv_tiled: the SWIGLU pair is column-parallel: nothing to hand off");
    if (h->send_seq)
      for (int r = 0; r < h->world; ++r)
        B200_CHECK_ARG(h->peer_y[r] && (reinterpret_cast<uintptr_t>(h->peer_y[r]) & 3) == 0, "b200_qgemv_tiled: tp producer needs a 4-byte aligned peer_y for every rank");
    if (h->slots)
      B200_CHECK_ARG(h->recv_seq && h->recv_done && h->slot_stride >= (int64_t)a->batch * a->K * 4 && h->slot_stride % 16 == 0 &&
                         (reinterpret_cast<uintptr_t>(h->slots) & 15) == 0,
                     "b200_qgemv_tiled: tp consumer needs recv_seq, recv_done and 16-byte aligned slots of >= batch * K * 4 bytes");
  }
  // more than one PLAIN row (no norm / delta prologue, no tensor-parallel hand-off): one row tile per warp (qgemv_nb.cu)
  if (a->batch >= g_nb_min_batch.load() && (g_stream_depth.load() == 0 || g_stream_depth.load() == 99)) {
    const bool ws_ok = dry || (a->workspace && a->workspace_bytes >= qgemv_v4_workspace_bytes() + qgemv_nb_workspace_bytes());
    if (ws_ok) {
      uint8_t* ws = dry ? nullptr : reinterpret_cast<uint8_t*>(a->workspace) + qgemv_v4_workspace_bytes();
      const int rc = qgemv_nb_launch(a, ws, dry ? nullptr : qgemv_v4_take_trace_slot(), (cudaStream_t)stream, dry, g_stream_depth.load() == 99 ? 1 : 0);
      if (rc != B200_ERR_UNSUPPORTED) return rc;
    }
  }
  // stream-K kernel (qgemv_v4.cu) unless forced off (tune depth 98) or the shape is outside its plan
  if (g_stream_depth.load() != 98) {
    const int rc = qgemv_v4_launch(a, g_stream_depth.load() == 99 ? 1 : g_stream_depth.load() == 97 ? 3 : 0, (cudaStream_t)stream, dry);  // 99: stream only, 97: prologue cycle stamps
    if (rc != B200_ERR_UNSUPPORTED) return rc;
  }
  if (fold > 1) {
    set_error("b200_qgemv_tiled: batch %d x K %lld (fold %d) does not fit the stream-K kernel's shared memory plan", a->batch, (long long)a->K, fold);
    return B200_ERR_UNSUPPORTED;
  }
  if (a->attn) {
    set_error("b200_qgemv_tiled: the fused attention step needs batch 1, n_head == n_query_groups, head size 128, N = 3 * n_head * 128 >= 16 * SMs "
              "and the STORE epilogue; call b200_attn_decode separately");
    return B200_ERR_UNSUPPORTED;
  }

  StreamParams p{};
  const bool swiglu = a->epilogue == B200_EPI_SWIGLU;
  p.tA = reinterpret_cast<const uint8_t*>(a->tiled);
  p.tB = swiglu ? reinterpret_cast<const uint8_t*>(a->tiled2) : p.tA;
  p.N = a->N; p.K = a->K; p.nkb = (int)(a->K / 64);
  p.ntiles = (int)((a->N + 15) / 16);
  p.nunits = p.ntiles;
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
  p.err = (a->workspace && a->workspace_bytes >= 4) ? reinterpret_cast<uint32_t*>(a->workspace) : nullptr;
  p.debug = g_stream_depth.load() == 99 ? 1 : 0;
  if (g_stream_pf.load() && a->prefetch && a->prefetch_bytes >= 16) { p.pf = reinterpret_cast<const uint8_t*>(a->prefetch); p.pf_bytes = a->prefetch_bytes; }

  if (a->tp) {
    const b200_tp_handoff* h = a->tp;
    B200_CHECK_ARG(h->world >= 2 && h->world <= 8 && h->rank >= 0 && h->rank < h->world, "b200_qgemv_tiled: tp world %d / rank %d out of range", h->world, h->rank);
    B200_CHECK_ARG(!(h->slots && a->delta), "b200_qgemv_tiled: tp slots replace delta (pass delta = NULL)");
    B200_CHECK_ARG(!(h->send_seq && swiglu), "b200_qgemv_tiled: the SWIGLU pair is column-parallel: nothing to hand off");
    p.tp_world = h->world;
    if (h->send_seq) {
      for (int r = 0; r < h->world; ++r) {
        B200_CHECK_ARG(h->peer_y[r] && (reinterpret_cast<uintptr_t>(h->peer_y[r]) & 3) == 0, "b200_qgemv_tiled: tp producer needs a 4-byte aligned peer_y for every rank");
        p.tp_peer_y[r] = reinterpret_cast<uint32_t*>(h->peer_y[r]);
      }
      p.tp_send_seq = reinterpret_cast<const uint32_t*>(h->send_seq);
    }
    if (h->slots) {
      B200_CHECK_ARG(h->recv_seq && h->recv_done && h->slot_stride >= (int64_t)a->batch * a->K * 4 && h->slot_stride % 16 == 0 &&
                         (reinterpret_cast<uintptr_t>(h->slots) & 15) == 0,
                     "b200_qgemv_tiled: tp consumer needs recv_seq, recv_done and 16-byte aligned slots of >= batch * K * 4 bytes");
      p.tp_slots = reinterpret_cast<const uint32_t*>(h->slots);
      p.tp_slot_stride = h->slot_stride / 4;
      p.tp_recv_seq = reinterpret_cast<uint32_t*>(h->recv_seq);
      p.tp_recv_done = reinterpret_cast<uint32_t*>(h->recv_done);
    }
  }

  const int NB = a->batch > 8 ? 2 : 1;
  const int COLS = 8 * NB;
  // mirror of the kernel's shared-memory carve-up (assuming the dynamic window starts within the first 1 KB)
  const size_t stage = 2 * (size_t)kCW * (int4 ? 1088 : 576);
  auto ring_slots = [&](int red_bufs) -> size_t {
    size_t off = 1024 + 128 + 128;
    off += (size_t)red_bufs * kCW * 2 * 16 * COLS * 4 + kCW * 16 * 4 + 16 * 4;
    if (int4) off += ((((size_t)a->batch * p.nkb) + 3) & ~(size_t)3) * 4;
    off += (((size_t)a->batch * p.xstride * 2) + 127) & ~(size_t)127;
    const size_t end = 1024 + (size_t)kDynSmem;
    if (int4) return off < end ? (end - off) / stage : 0;   // no table
    const size_t lut_at = (off + 0xFFFF) & ~(size_t)0xFFFF, ring_c = lut_at + 0x10000;
    if (ring_c > end) return 0;
    return (lut_at - off) / stage + (end - ring_c) / stage;
  };
  p.red_bufs = ring_slots(2) >= 3 ? 2 : 1;
  if (ring_slots(p.red_bufs) < 2) {
    set_error("b200_qgemv_tiled: batch %d x K %lld does not fit the kernel's shared memory plan (activation rows need %zu bytes)",
              a->batch, (long long)a->K, (size_t)a->batch * p.xstride * 2);
    return B200_ERR_UNSUPPORTED;
  }
  if (dry) return B200_OK;
  const int steps = swiglu ? p.nunits : (p.nunits + 1) / 2;
  int gx = sm_count();
  if (gx > steps) gx = steps;
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e;
  if (int4) e = NB == 2 ? launch_stream<2, true>(p, gx, st) : launch_stream<1, true>(p, gx, st);
  else e = NB == 2 ? launch_stream<2, false>(p, gx, st) : launch_stream<1, false>(p, gx, st);
  if (e != cudaSuccess) {
    set_error("b200_qgemv_tiled: launch failed: %s", cudaGetErrorString(e));
    return B200_ERR_CUDA;
  }
  return B200_OK;
}

int b200_qgemv_tiled(const b200_qgemv_tiled_args* a, b200_stream_t stream) { return qgemv_tiled_impl(a, stream, false); }

int b200_qgemv_tiled_error(void* workspace, b200_stream_t stream, unsigned int* code) {
  B200_CHECK_ARG(workspace && code, "b200_qgemv_tiled_error: null argument");
  B200_CUDA(cudaMemcpyAsync(code, workspace, 4, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  B200_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  if (*code) B200_CUDA(cudaMemsetAsync(workspace, 0, 4, (cudaStream_t)stream));
  return B200_OK;
}

int b200_qgemv_tiled_supported(const b200_qgemv_tiled_args* a) { return qgemv_tiled_impl(a, nullptr, true) == B200_OK ? 1 : 0; }

}  // extern "C"
