// Batched dequant-GEMV (2..16 activation rows) over the tiled 4-bit layout: ONE 16-ROW TILE PER CONSUMER WARP.
//
// Replaces, for more than one row, the same reference code as qgemv_v4.cu: nn.Linear.forward on a quantized weight with a handful
// of rows (lit_gpt/model.py:619 attn, :656 proj, :713-716 fc_1 / fc_2 / SwiGLU / proj, :519 lm_head) under
// `--quantize bnb.nf4 / bnb.fp4 / gptq.int4` (quantize/bnb.py Linear4bit, quantize/gptq.py ColBlockQuantizedLinear).
//
// Why a second kernel.  qgemv_v4.cu splits K over the 16 consumer warps of a CTA: right for one row (the activation row lives in
// registers / 8 KB of shared memory), wrong for sixteen: every CTA needs ALL rows (16 x K bf16 = 128 - 442 KB: k-windows, a
// staging pass per window), every (tile, window) ends in a 16 KB cross-warp reduction through shared memory and one reducer
// warp, and a warp sees two records between hand-offs.  Measured at 16 rows: 0.13 of the HBM roofline.  Here
//   * a GROUP is 16 consecutive row tiles, one per consumer warp; a STEP is one k-unit (RU records, 256 k) of all 16 tiles;
//   * the ring stage of a step holds the 16 weight units (16 bulk copies, 2.2 KB contiguous each) AND the step's slice of the
//     activation rows (B bulk copies of 512 B): the rows are STREAMED with the weights, never staged, whatever K is;
//   * a warp keeps its tile's 16 x 16 partial sums in registers for the whole group: no shared-memory reduction, no reducer warp;
//   * stream-K over steps: CTA c owns steps [S c / G, S (c + 1) / G).  A group cut by a range boundary is summed in CTA order by
//     the CTA holding its first step (per-warp flags + 1 KB partials in the workspace, bounded waits): deterministic.
// int4: a = bf16(128 + code) - bf16(128 + zero) is exact, so the zero point is folded into the A fragment (one HADD2.BF16 per
// register) and no per-block activation sums are needed.
// Plain rows only: RMSNorm / residual-delta prologues and the tensor-parallel hand-off stay with qgemv_v4.cu (the caller runs
// b200_rmsnorm as its own launch above 8 rows, lit_gpt_b200/model.py).
#include "qgemv_v4.cuh"
#include "stream_common.cuh"

namespace b200 {

namespace {

constexpr int kNbCW = 16;                       // consumer warps = tiles per group
constexpr int kNbThreads = (kNbCW + 1) * 32;    // + producer warp
constexpr int kNbSmem = 227 * 1024;
constexpr int kNbMaxCtas = 160;
constexpr int kNbMaxStages = 8;
constexpr size_t kNbWsPerCta = 16384;           // [warp][2][32] x 16 B: four fp32 partial sums each, bit 0 = valid

struct NbParams {
  const uint8_t* tA;
  const uint8_t* tB;
  int64_t N;
  int K, upt, ts, ntf, steps;       // units per tile; tiles per entity (SwiGLU: 2); flat tiles; ngroups * upt
  int qtype, batch, epilogue, debug;   // debug 1: stream only (development: the ring without the arithmetic)
  const __nv_bfloat16* x;
  const __nv_bfloat16* bias;
  const __nv_bfloat16* bias2;
  const __nv_bfloat16* residual;
  __nv_bfloat16* y;
  uint32_t* err;                    // word 0 of the caller's workspace (b200_qgemv_tiled_error)
  uint8_t* ws;                      // this kernel's region: the partial-sum packets
  unsigned long long* trace;
};

struct NbPlan {
  uint32_t bars, xch, ringA, lut, ringC;
  int nA, nC, S;
};

// shared-memory carve-up by absolute shared-window address (nf4 / fp4: the table sits at a 64 KB aligned address, see qgemv_v4.cu)
__host__ __device__ inline NbPlan nb_plan(uint32_t b0, int stage, bool int4) {
  NbPlan pl{};
  uint32_t off = (b0 + 127u) & ~127u;
  pl.bars = off; off += 256;
  const uint32_t end = b0 + (uint32_t)kNbSmem;
  if (int4) {
    pl.xch = off; off += 16384;      // SwiGLU exchange: 2 slots x 8 warp pairs x 8 x 128 B
    pl.ringA = off; pl.lut = 0; pl.ringC = end;
    pl.nA = (int)((end - off) / stage); pl.nC = 0;
  } else {
    pl.ringA = off; pl.xch = 0;      // the exchange lives in the upper halves of the 256-byte table rows
    int best = -1;
    for (uint32_t cand = 0x10000u; cand + 0x10000u <= end; cand += 0x10000u) {
      if (cand < off) continue;
      const int nA = (int)((cand - off) / stage), nC = (int)((end - cand - 0x10000u) / stage);
      if (nA + nC > best) { best = nA + nC; pl.lut = cand; pl.ringC = cand + 0x10000u; pl.nA = nA; pl.nC = nC; }
    }
    if (best < 0) { pl.lut = 0; pl.nA = pl.nC = 0; }
  }
  pl.S = pl.nA + pl.nC < kNbMaxStages ? pl.nA + pl.nC : kNbMaxStages;
  return pl;
}

__device__ __forceinline__ void nb_stamp(const NbParams& p, int ev) {
  if (p.trace) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    p.trace[(size_t)blockIdx.x * 8 + ev] = t;
  }
}

__device__ __forceinline__ uint32_t bf16x2_sub(uint32_t a, uint32_t b) {
  const __nv_bfloat162 r = __hsub2(*reinterpret_cast<const __nv_bfloat162*>(&a), *reinterpret_cast<const __nv_bfloat162*>(&b));
  return *reinterpret_cast<const uint32_t*>(&r);
}

// NB = 1: batch <= 8 (one n8 accumulator tile), NB = 2: batch <= 16
template <int RU, bool INT4, int NB>
__global__ void __launch_bounds__(kNbThreads, 1) qgemv_nb_kernel(const NbParams p) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  constexpr int REC = INT4 ? 1088 : 576;
  constexpr int BPR = INT4 ? 2 : 1;             // 64-wide k-blocks per record
  constexpr int UB = RU * REC;                  // unit bytes
  constexpr int NBLK = RU * BPR;                // k-blocks per unit
  constexpr int UK = NBLK * 64;                 // k per step
  constexpr int XROW = UK * 2 + 16;             // activation slice row pitch: + 16 B keeps the fragment loads conflict-free
  constexpr int XROWS = 8 * NB;
  constexpr int STAGE = (kNbCW * UB + XROWS * XROW + 127) & ~127;
  constexpr int GU = 8;                         // contributors gathered per L2 round trip

  const uint32_t b0 = smem_u32(smem_raw);
  uint8_t* const sb = smem_raw - b0;            // generic pointer to shared-window address 0
  const NbPlan pl = nb_plan(b0, STAGE, INT4);
  const int S = pl.S;
  if (S < 2) __trap();                          // the host sized the launch so that this cannot happen
  const uint32_t full_a = pl.bars, empty_a = pl.bars + 64;
  auto stage_addr = [&](int s) -> uint32_t { return s < pl.nA ? pl.ringA + s * STAGE : pl.ringC + (s - pl.nA) * STAGE; };

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int G = gridDim.x, cta = blockIdx.x;
  const int B = p.batch, upt = p.upt, TS = p.ts;
  auto range_start = [&](int c) -> int { return (int)(((int64_t)p.steps * c) / G); };
  const int s0 = range_start(cta), s1 = range_start(cta + 1);

  if (tid == 0) {
    nb_stamp(p, 0);
    for (int s = 0; s < kNbMaxStages; ++s) { mbar_init(full_a + s * 8, 1); mbar_init(empty_a + s * 8, kNbCW); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  griddep_launch();

  if (warp == kNbCW) {
    // ================================ PRODUCER WARP ==============================================================
    // lane l < 16 copies tile l's unit, lanes 16.. copy the activation rows' slice; lane 0 arms the barrier first
    const uint64_t pol = make_evict_first_policy();
    auto issue_w = [&](int st, int slot) {
      const int gi = st / upt, ku = st - gi * upt;
      const int tf = gi * kNbCW + lane;
      if (lane < kNbCW && tf < p.ntf) {
        const uint8_t* src = (TS == 2 && (tf & 1)) ? p.tB : p.tA;
        const int mt = TS == 2 ? (tf >> 1) : tf;
        bulk_g2s(stage_addr(slot) + (uint32_t)lane * UB, src + ((int64_t)mt * upt + ku) * UB, (uint32_t)UB, full_a + slot * 8, pol);
      }
    };
    auto issue_x = [&](int st, int slot) {
      const int ku = st % upt, b = lane - 16;
      if (b >= 0 && b < B)
        bulk_g2s_plain(stage_addr(slot) + (uint32_t)(kNbCW * UB) + (uint32_t)b * XROW, p.x + (int64_t)b * p.K + (int64_t)ku * UK, (uint32_t)(UK * 2),
                       full_a + slot * 8);
    };
    auto arm = [&](int st, int slot) {
      if (lane == 0) {
        const int gi = st / upt;
        const int nt = min(kNbCW, p.ntf - gi * kNbCW);
        mbar_expect_tx(full_a + slot * 8, (uint32_t)nt * UB + (uint32_t)B * (UK * 2));
      }
      __syncwarp();
    };
    const int n = s1 - s0;
    const int pre = n < S ? n : S;
    // the weights do not depend on the previous launch: the first S steps' units are in flight before the dependency resolves
    for (int i = 0; i < pre; ++i) { arm(s0 + i, i); issue_w(s0 + i, i); }
    griddep_wait();
    if (lane == 0) nb_stamp(p, 1);
    for (int i = 0; i < pre; ++i) issue_x(s0 + i, i);
    int slot = 0;
    uint32_t pph = 0;              // parity of the EMPTY phase to wait for (fill f waits for the consumers' release of fill f - 1)
    for (int i = pre; i < n; ++i) {
      if (lane == 0) mbar_wait(empty_a + slot * 8, pph);
      __syncwarp();
      arm(s0 + i, slot);
      issue_w(s0 + i, slot);
      issue_x(s0 + i, slot);
      if (++slot == S) { slot = 0; pph ^= 1u; }
    }
    if (lane == 0) nb_stamp(p, 6);
    return;
  }

  // ================================== CONSUMER WARPS ===============================================================
  const int g = lane >> 2, t = lane & 3;
  const uint32_t lut_a = pl.lut;
  if (!INT4) {
    // table entry e = warp + 16 k, lane slot `lane` at byte e * 256 + lane * 4: element 0 (low half) = code[k], element 1 = code[warp]
    uint32_t* lut = reinterpret_cast<uint32_t*>(sb + lut_a);
    const float* cb = p.qtype == B200_Q_NF4 ? kNF4 : kFP4;
    const float mine = cb[warp];
#pragma unroll
    for (int k = 0; k < 16; ++k) lut[(warp + 16 * k) * 64 + lane] = pack_bf16(cb[k], mine);
    asm volatile("bar.sync 1, %0;" ::"n"(kNbCW * 32) : "memory");
  }
  if (tid == 0) nb_stamp(p, 2);
  const uint32_t lane_full = lut_a | ((uint32_t)lane << 2);
  auto LUT = [&](uint32_t word, uint32_t sel) -> uint32_t {
    uint32_t v;
    asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(__byte_perm(word, lane_full, sel)));
    return v;
  };
  uint32_t i4mask = 0x000F000Fu, i4magic = 0x43004300u;
  asm volatile("" : "+r"(i4mask), "+r"(i4magic));

  // activation fragments: lane (g, t) feeds batch column g (+ 8) with k = 16 t .. 16 t + 15 of the k-block
  const uint32_t xo0 = (uint32_t)(kNbCW * UB) + (uint32_t)((g < B ? g : 0) * XROW) + (uint32_t)t * 32u;
  const uint32_t xo1 = (uint32_t)(kNbCW * UB) + (uint32_t)((g + 8 < B ? g + 8 : 0) * XROW) + (uint32_t)t * 32u;

  float acc[NB][4];
#pragma unroll
  for (int nb = 0; nb < NB; ++nb) {
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[nb][i] = 0.f;
  }

  // one 64-wide k-block of this warp's tile: w = this lane's 16 B {row g bytes, row g + 8 bytes}; sc = scale words of rows g, g + 8;
  // xa = shared address of the block's activation slice (row 0, k 0)
  auto compute = [&](const uint4 w, const uint2 sc, uint32_t xa) {
    uint32_t xb[NB][8];
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
      const uint32_t a = xa + (nb == 0 ? xo0 : xo1);
      asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(xb[nb][0]), "=r"(xb[nb][1]), "=r"(xb[nb][2]), "=r"(xb[nb][3]) : "r"(a));
      asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(xb[nb][4]), "=r"(xb[nb][5]), "=r"(xb[nb][6]), "=r"(xb[nb][7]) : "r"(a + 16u));
    }
    float pr[NB][4];
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
#pragma unroll
      for (int i = 0; i < 4; ++i) pr[nb][i] = 0.f;
    }
    float sa, sbv;
    uint32_t zpa = 0, zpb = 0;
    if constexpr (INT4) {
      const __half2 ha = *reinterpret_cast<const __half2*>(&sc.x), hb = *reinterpret_cast<const __half2*>(&sc.y);
      sa = __low2float(ha); sbv = __low2float(hb);
      const float za = __high2float(ha) + 128.f, zb = __high2float(hb) + 128.f;   // integers <= 144: exact in bf16
      zpa = pack_bf16(za, za); zpb = pack_bf16(zb, zb);
    } else {
      sa = __uint_as_float(sc.x); sbv = __uint_as_float(sc.y);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t wa = (j < 2) ? w.x : w.y, wb = (j < 2) ? w.z : w.w;
      uint32_t a0, a1, a2, a3;
      if constexpr (INT4) {
        // even codes in the low, odd codes in the high half-word: one shift + one LOP3 yields the bf16 pair (128 + e_2p, 128 + e_2p+1)
        const int sh = 8 * (j & 1);
        auto ex = [&](uint32_t v, int s) -> uint32_t {
          uint32_t r;
          asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(r) : "r"(s ? v >> s : v), "r"(i4mask), "r"(i4magic));
          return r;
        };
        a0 = bf16x2_sub(ex(wa, sh), zpa); a1 = bf16x2_sub(ex(wb, sh), zpb);
        a2 = bf16x2_sub(ex(wa, sh + 4), zpa); a3 = bf16x2_sub(ex(wb, sh + 4), zpb);
      } else {
        const uint32_t sel0 = (j & 1) ? 0x7624u : 0x7604u, sel1 = (j & 1) ? 0x7634u : 0x7614u;
        a0 = LUT(wa, sel0); a2 = LUT(wa, sel1); a1 = LUT(wb, sel0); a3 = LUT(wb, sel1);
      }
#pragma unroll
      for (int nb = 0; nb < NB; ++nb) mma16816(pr[nb], a0, a1, a2, a3, xb[nb][2 * j], xb[nb][2 * j + 1]);
    }
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
      acc[nb][0] = fmaf(sa, pr[nb][0], acc[nb][0]);
      acc[nb][1] = fmaf(sa, pr[nb][1], acc[nb][1]);
      acc[nb][2] = fmaf(sbv, pr[nb][2], acc[nb][2]);
      acc[nb][3] = fmaf(sbv, pr[nb][3], acc[nb][3]);
    }
  };

  // partial sums of a group cut by a range boundary travel as self-validating words: fp32 with the mantissa's last bit forced to
  // 1 (a 2^-24 relative change, far below the bf16 rounding of the output; the workspace rests at zero = "not there yet").  The
  // owner polls the data itself -- one L2 round trip, no fence + flag on the contributor's side -- and writes zeros back.
  uint4* const parts = reinterpret_cast<uint4*>(p.ws);
  // SwiGLU exchange slot (epilogue parity, warp pair, value j): one 128-byte row per (slot, pair, j)
  auto xch_at = [&](int par, int pair, int j) -> float* {
    const uint32_t r = (uint32_t)((par * 8 + pair) * 8 + j);
    return reinterpret_cast<float*>(sb + (INT4 ? pl.xch + r * 128u : lut_a + 128u + r * 256u)) + lane;
  };

  griddep_wait();   // complete long before the first stage arrives (the producer issues the activation slices after its own wait)
  int slot = 0, epi = 0;
  uint32_t ph = 0;
  int st = s0;
  while (st < s1) {
    const int gi = st / upt, g_begin = gi * upt, g_end = g_begin + upt;
    const int seg_end = s1 < g_end ? s1 : g_end;
    const bool contrib = st > g_begin;      // the group began in a lower CTA: this CTA contributes partial sums
    const bool cut = seg_end < g_end;       // the group continues in higher CTAs
    const int tf = gi * kNbCW + warp;       // this warp's flat tile
    const bool active = tf < p.ntf;
    if (active && !contrib && p.epilogue == B200_EPI_RESIDUAL) {
      // this warp will finish the tile: fetch the residual values its epilogue adds NOW, asynchronously and without registers
      // (cp.async of the aligned 4-byte word holding each bf16, into the exchange rows -- SwiGLU and RESIDUAL exclude each other).
      // In the tail every dependent L2 round trip costs 2 - 3 us (the next launch's CTAs are already streaming on the SMs the
      // contributors left), so the epilogue must not start one
#pragma unroll
      for (int j = 0; j < NB * 4; ++j) {
        const int64_t n = (int64_t)tf * 16 + g + 8 * ((j & 3) >> 1);
        const int c = 8 * (j >> 2) + 2 * t + (j & 1);
        if (c < B && n < p.N) {
          const uint32_t dst = smem_u32(xch_at(0, 0, warp * 8 + j));
          const __nv_bfloat16* src = p.residual + (int64_t)c * p.N + n;
          asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(reinterpret_cast<uintptr_t>(src) & ~uintptr_t(3)) : "memory");
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    }
    for (; st < seg_end; ++st) {
      mbar_wait(full_a + slot * 8, ph);
      if (tid == 0 && st == s0) nb_stamp(p, 3);
      if (active && p.debug != 1) {
        const uint32_t sbase = stage_addr(slot);
        const uint32_t sp = sbase + (uint32_t)warp * UB;
        uint4 wv[RU][BPR];
        uint2 sc[RU];
#pragma unroll
        for (int r = 0; r < RU; ++r) {
#pragma unroll
          for (int b = 0; b < BPR; ++b) {
            const uint32_t wa = sp + r * REC + b * 512 + lane * 16;
            asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(wv[r][b].x), "=r"(wv[r][b].y), "=r"(wv[r][b].z), "=r"(wv[r][b].w) : "r"(wa));
          }
          asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(sc[r].x), "=r"(sc[r].y) : "r"(sp + r * REC + BPR * 512 + g * 8));
        }
#pragma unroll
        for (int r = 0; r < RU; ++r) {
#pragma unroll
          for (int b = 0; b < BPR; ++b) compute(wv[r][b], sc[r], sbase + (uint32_t)(r * BPR + b) * 128u);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(empty_a + slot * 8);
      if (++slot == S) { slot = 0; ph ^= 1u; }
    }
    if (active) {
      if (contrib) {
        // CONTRIBUTOR: publish this warp's partial sums; the owner's warp of the same index (same lanes) adds them
        uint4* wd = parts + ((size_t)cta * kNbCW + warp) * 64;
#pragma unroll
        for (int nb = 0; nb < NB; ++nb)
          asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(wd + nb * 32 + lane), "r"(__float_as_uint(acc[nb][0]) | 1u),
                       "r"(__float_as_uint(acc[nb][1]) | 1u), "r"(__float_as_uint(acc[nb][2]) | 1u), "r"(__float_as_uint(acc[nb][3]) | 1u) : "memory");
      } else {
        // OWNER of this warp's tile (its residual values were copied to shared memory when the segment began)
        const bool swi = p.epilogue == B200_EPI_SWIGLU;
        if (cut) {
          if (tid == 0) nb_stamp(p, 7);
          // the group continues in higher CTAs: add their partial sums in CTA order (deterministic), GU contributors per round trip
          const int c_lo = cta + 1;
          int c_hi = (int)(((int64_t)g_end * G + p.steps - 1) / p.steps);   // first CTA whose range starts at or after the group's end
          if (c_hi > G) c_hi = G;
          for (int cb = c_lo; cb < c_hi; cb += GU) {
            uint4 pk[GU][NB];
            const long long t0 = clock64();
            bool all;
            do {
              all = true;
#pragma unroll
              for (int u = 0; u < GU; ++u) {
                const uint4* wd = parts + ((size_t)(cb + u) * kNbCW + warp) * 64;
#pragma unroll
                for (int nb = 0; nb < NB; ++nb) {
                  pk[u][nb] = make_uint4(1u, 1u, 1u, 1u);   // (+1.4e-45: nothing)
                  if (cb + u < c_hi)
                    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(pk[u][nb].x), "=r"(pk[u][nb].y), "=r"(pk[u][nb].z), "=r"(pk[u][nb].w)
                                 : "l"(wd + nb * 32 + lane) : "memory");
                }
              }
#pragma unroll
              for (int u = 0; u < GU; ++u) {
#pragma unroll
                for (int nb = 0; nb < NB; ++nb) all = all && ((pk[u][nb].x & pk[u][nb].y & pk[u][nb].z & pk[u][nb].w) & 1u);
              }
              if (!all && clock64() - t0 > 4000000000LL) { *reinterpret_cast<volatile uint32_t*>(p.err) = 1u; break; }  // error word, no hang
            } while (!all);
#pragma unroll
            for (int u = 0; u < GU; ++u) {
              uint4* wd = parts + ((size_t)(cb + u) * kNbCW + warp) * 64;
#pragma unroll
              for (int nb = 0; nb < NB; ++nb) {
                if (cb + u < c_hi) {
                  acc[nb][0] += __uint_as_float(pk[u][nb].x); acc[nb][1] += __uint_as_float(pk[u][nb].y);
                  acc[nb][2] += __uint_as_float(pk[u][nb].z); acc[nb][3] += __uint_as_float(pk[u][nb].w);
                  // self-cleaning: ready for the next launch
                  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %1, %1, %1};" ::"l"(wd + nb * 32 + lane), "r"(0u) : "memory");
                }
              }
            }
          }
        }
        // ---- epilogue of this warp's tile: lane holds rows g / g + 8, columns 8 nb + 2 t + {0, 1} -------------------------------
        if (swi) {
          const int mt = tf >> 1, pair = warp >> 1;
          const bool up = tf & 1;
          const __nv_bfloat16* bs = up ? p.bias2 : p.bias;
          float v[NB][4];
#pragma unroll
          for (int nb = 0; nb < NB; ++nb) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const int64_t n = (int64_t)mt * 16 + g + 8 * (i >> 1);
              v[nb][i] = round_bf16(acc[nb][i] + ((bs && n < p.N) ? __bfloat162float(bs[n]) : 0.f));
            }
          }
          if (up) {
#pragma unroll
            for (int nb = 0; nb < NB; ++nb) {
#pragma unroll
              for (int i = 0; i < 4; ++i) *xch_at(epi & 1, pair, nb * 4 + i) = v[nb][i];
            }
          }
          asm volatile("bar.sync %0, 64;" ::"r"(2 + pair) : "memory");
          if (!up) {
#pragma unroll
            for (int nb = 0; nb < NB; ++nb) {
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const int64_t n = (int64_t)mt * 16 + g + 8 * (i >> 1);
                const int c = 8 * nb + 2 * t + (i & 1);
                const float b = *xch_at(epi & 1, pair, nb * 4 + i);
                if (c < B && n < p.N) p.y[(int64_t)c * p.N + n] = __float2bfloat16_rn(silu_bf16_s(v[nb][i]) * b);
              }
            }
          }
          ++epi;
        } else {
          int at[NB][4];                 // element offsets into y / residual (the host keeps batch * N below 2^31)
          float add[NB][4], res[NB][4];
#pragma unroll
          for (int nb = 0; nb < NB; ++nb) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const int64_t n = (int64_t)tf * 16 + g + 8 * (i >> 1);
              const int c = 8 * nb + 2 * t + (i & 1);
              const bool ok = c < B && n < p.N;
              at[nb][i] = ok ? (int)(c * p.N + n) : -1;
              add[nb][i] = (ok && p.bias) ? __bfloat162float(p.bias[n]) : 0.f;
            }
          }
          if (p.epilogue == B200_EPI_RESIDUAL) asm volatile("cp.async.wait_all;" ::: "memory");
#pragma unroll
          for (int nb = 0; nb < NB; ++nb) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              res[nb][i] = 0.f;
              if (at[nb][i] >= 0 && p.epilogue == B200_EPI_RESIDUAL) {
                const uint32_t wd = *reinterpret_cast<const volatile uint32_t*>(xch_at(0, 0, warp * 8 + nb * 4 + i));
                res[nb][i] = bf16_bits_to_float((at[nb][i] & 1) ? (wd >> 16) : (wd & 0xffffu));   // y / residual are 4-byte aligned rows (N even)
              }
            }
          }
#pragma unroll
          for (int nb = 0; nb < NB; ++nb) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              if (at[nb][i] < 0) continue;
              float h = acc[nb][i] + add[nb][i];
              if (p.epilogue == B200_EPI_RESIDUAL) h = round_bf16(h) + res[nb][i];
              p.y[at[nb][i]] = __float2bfloat16_rn(h);
            }
          }
        }
      }
    }
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
#pragma unroll
      for (int i = 0; i < 4; ++i) acc[nb][i] = 0.f;
    }
    if (tid == 0 && st >= s1) nb_stamp(p, 4);
  }
  if (tid == 0) nb_stamp(p, 5);
}

template <int RU, bool INT4, int NB>
cudaError_t launch_nb(const NbParams& p, int gx, cudaStream_t st) {
  static PerDeviceOnce once;  // per instantiation
  if (once.need()) {
    cudaError_t e = cudaFuncSetAttribute(qgemv_nb_kernel<RU, INT4, NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, kNbSmem);
    if (e != cudaSuccess) return e;
    once.done();
  }
  return launch(qgemv_nb_kernel<RU, INT4, NB>, dim3(gx), dim3(kNbThreads), (size_t)kNbSmem, st, p);
}

template <bool INT4, int NB>
cudaError_t launch_nb_ru(int ru, const NbParams& p, int gx, cudaStream_t st) {
  if (INT4) return ru == 2 ? launch_nb<2, INT4, NB>(p, gx, st) : launch_nb<1, INT4, NB>(p, gx, st);
  return ru == 4 ? launch_nb<4, INT4, NB>(p, gx, st) : ru == 2 ? launch_nb<2, INT4, NB>(p, gx, st) : launch_nb<1, INT4, NB>(p, gx, st);
}

}  // namespace

static std::atomic<int> g_nb_ru_max{4};   // development: largest unit (records per step and tile)
void qgemv_nb_tune_ru(int ru_max) { g_nb_ru_max = ru_max > 0 ? ru_max : 4; }

size_t qgemv_nb_workspace_bytes() { return (size_t)kNbMaxCtas * kNbWsPerCta; }

int qgemv_nb_launch(const b200_qgemv_tiled_args* a, uint8_t* ws, unsigned long long* trace, cudaStream_t st, bool dry, int debug) {
  if (a->batch < 2 || a->batch > 16 || a->fold > 1 || a->norm_weight || a->delta || a->x_out || a->tp || a->attn) return B200_ERR_UNSUPPORTED;
  const bool int4 = a->qtype == B200_Q_INT4;
  const bool swiglu = a->epilogue == B200_EPI_SWIGLU;
  const int rec_k = int4 ? 128 : 64;
  if (a->K % rec_k || a->K > 0x3fffffff || a->N * 16 > 0x7fffffff || (a->N & 1)) return B200_ERR_UNSUPPORTED;
  if (!dry && a->epilogue == B200_EPI_RESIDUAL && (reinterpret_cast<uintptr_t>(a->residual) & 3)) return B200_ERR_UNSUPPORTED;   // 4-byte words of residual are copied
  const int nrec = (int)(a->K / rec_k);
  int ru = int4 ? (nrec % 2 == 0 ? 2 : 1) : (nrec % 4 == 0 ? 4 : nrec % 2 == 0 ? 2 : 1);
  while (ru > g_nb_ru_max.load()) ru >>= 1;
  const int nb = a->batch <= 8 ? 1 : 2;
  const int ub = ru * (int4 ? 1088 : 576), uk = ru * rec_k;
  const int stage = (kNbCW * ub + 8 * nb * (uk * 2 + 16) + 127) & ~127;
  if (nb_plan(1024u, stage, int4).S < 2) return B200_ERR_UNSUPPORTED;
  const int ts = swiglu ? 2 : 1;
  const int64_t ntf = ((a->N + 15) / 16) * ts;
  const int64_t ngroups = (ntf + kNbCW - 1) / kNbCW;
  const int upt = nrec / ru;
  const int64_t steps = ngroups * upt;
  if (ntf > 0x3fffffff || steps > 0x3fffffff) return B200_ERR_UNSUPPORTED;
  if (dry) return B200_OK;
  if (!ws) return B200_ERR_UNSUPPORTED;

  NbParams p{};
  p.tA = reinterpret_cast<const uint8_t*>(a->tiled);
  p.tB = swiglu ? reinterpret_cast<const uint8_t*>(a->tiled2) : p.tA;
  p.N = a->N; p.K = (int)a->K; p.upt = upt; p.ts = ts; p.ntf = (int)ntf; p.steps = (int)steps;
  p.qtype = a->qtype; p.batch = a->batch; p.epilogue = a->epilogue;
  p.x = reinterpret_cast<const __nv_bfloat16*>(a->x);
  p.bias = reinterpret_cast<const __nv_bfloat16*>(a->bias);
  p.bias2 = reinterpret_cast<const __nv_bfloat16*>(a->bias2);
  p.residual = reinterpret_cast<const __nv_bfloat16*>(a->residual);
  p.y = reinterpret_cast<__nv_bfloat16*>(a->y);
  p.err = reinterpret_cast<uint32_t*>(a->workspace);
  p.ws = ws;
  p.trace = trace;
  p.debug = debug;
  int gx = sm_count();
  if (gx > kNbMaxCtas) gx = kNbMaxCtas;
  if ((int64_t)gx > steps) gx = (int)steps;     // every CTA owns at least one step (the owners' waits rely on it)
  cudaError_t e;
  if (int4) e = nb == 1 ? launch_nb_ru<true, 1>(ru, p, gx, st) : launch_nb_ru<true, 2>(ru, p, gx, st);
  else e = nb == 1 ? launch_nb_ru<false, 1>(ru, p, gx, st) : launch_nb_ru<false, 2>(ru, p, gx, st);
  if (e != cudaSuccess) {
    set_error("b200_qgemv_tiled: launch failed: %s", cudaGetErrorString(e));
    return B200_ERR_CUDA;
  }
  return B200_OK;
}

}  // namespace b200
