// Stream-K fused dequant-GEMV over the TILED weight layout (csrc/tile.cu), batch 1..16.
//
// Replaces bnb.nn.Linear4bit.forward at lit_gpt/model.py:619,656,713,714,716,519.
//
// Work decomposition.  The tiled matrix is ONE contiguous byte stream of records (16 rows x 64 k, nf4: 576 B; int4:
// 16 rows x 128 k, 1088 B), tile-major.  A UNIT is RU consecutive records of one tile (nf4 RU = 4: 256 k, 2304 B).  An
// ENTITY is a tile (SwiGLU: a gate tile + the up tile of the same rows).  Every CTA takes a contiguous range of units:
// whole entities when that balances well enough, else equal unit counts (stream-K: every SM streams the same number
// of bytes +- one unit whatever N and K are).  An entity cut by a range boundary is finished by the CTA holding its
// first unit (the OWNER); the other CTAs holding parts of it publish their fp32 partial sums to the workspace right
// after their first units (they meet the cut entity at the START of their range, the owner at the END of its own, so
// the hand-off latency is hidden) and the owner adds them in CTA order: deterministic, no atomics on data.
//
// Inside a CTA (576 threads):
//   warp 16  PRODUCER: one lane streams the range in stages of 16 units (nf4: 36,864 B, contiguous) with cp.async.bulk
//            into a shared-memory ring guarded by FULL / EMPTY mbarriers.
//   warps 0..15 CONSUMERS: warp w takes unit w of every stage.  Dequant: nf4 / fp4 one PRMT + one conflict-free LDS per
//            packed byte from a per-lane replicated 256-entry byte -> bf16x2 table at a 64 KB aligned shared address;
//            int4 one shift + one LOP3 per two weights.  MACs: mma.sync m16n8k16 against activation fragments (kept in
//            registers for the whole launch when a warp always meets the same k-range, else read from shared memory);
//            absmax / {scale, zero} applied in fp32 per 64 / 128 k.  When a warp leaves a tile it drops its partial sums
//            into a reduction buffer and arrives on that buffer's mbarrier: consumers never wait for each other (no
//            CTA-wide barrier in the loop).
//   warp 17  REDUCER: sums the 16 partials of a finished tile in fixed order, adds partner CTAs' partials for cut
//            entities, runs the epilogue (bias / residual / SwiGLU / tensor-parallel packets) and frees the buffer.
// Prologue (consumers): residual add + RMSNorm of the activation (lit_gpt/rmsnorm.py:19-25, model.py:591) into shared
// memory; for batch 1 the activation row lives in the unused halves of the 256 B table rows.
#include "qgemv_v4.cuh"
#include "stream_common.cuh"
#include "attn_body.cuh"

namespace b200 {

constexpr int kV4CW = 16;                     // consumer warps = units per stage
constexpr int kV4Threads = (kV4CW + 2) * 32;  // + producer + reducer
constexpr int kV4Consumers = kV4CW * 32;
constexpr int kV4MaxStages = 8;
constexpr int kV4DynSmem = 227 * 1024;
constexpr int kV4DynSmem2 = 110 * 1024;   // two co-resident CTAs per SM (experiment: consecutive launches overlap)
constexpr int kV4MaxCtas = 160;
constexpr size_t kV4WsData = 4096;            // workspace: [0] error word, [64 + 4 c] flag of CTA c, [4096 + 2048 c] partials of CTA c
constexpr size_t kV4WsPerCta = 2048;

struct V4Params {
  const uint8_t* tA;
  const uint8_t* tB;
  int64_t N, K;
  int nkb;          // K / 64
  int upt;          // units per tile
  int ts;           // tiles per entity: 1, SwiGLU 2 (gate tile, up tile)
  int U;            // total units
  int qtype, batch, epilogue, debug;
  const __nv_bfloat16* x;
  const __nv_bfloat16* delta;
  __nv_bfloat16* x_out;
  const __nv_bfloat16* norm_w;
  float eps;
  const __nv_bfloat16* bias;
  const __nv_bfloat16* bias2;
  const __nv_bfloat16* residual;
  __nv_bfloat16* y;
  int xstride;      // activation row stride in shared memory, bf16 elements
  int gap;          // 1: batch-1 activation row stored in the table gaps
  int split;        // 1: ranges cut entities (stream-K with partial hand-off); 0: every CTA owns whole entities
  uint8_t* ws;
  int tp_world;
  uint32_t* tp_peer_y[8];
  const uint32_t* tp_send_seq;
  const uint32_t* tp_slots;
  uint32_t* tp_reduced;      // two-stage exchange: local republish buffer (b200_tp_handoff::reduced), NULL = one stage
  int64_t tp_slot_stride;
  uint32_t* tp_recv_seq;
  uint32_t* tp_recv_done;
  const uint8_t* kvh_k; const uint8_t* kvh_v; const void* kvh_pos;   // K / V cache L2 hint (b200_kv_hint), kvh_k == NULL: none
  int kvh_i64, kvh_chunks, kvh_row, kvh_max; long long kvh_stride;
  // fused decode attention (b200_qgemv_tiled_args::attn): parameters, 128-thread sub-blocks per head, row tiles per head, counters
  AttnParams att; int att_nsplit, att_tph; unsigned int* att_cnt;
  int nbuf;         // batch > 1: reduction buffers (2, or 1 when that lets the activation rows fit without k-windows)
  int wu, nwin;     // k-windows (batch > 1 whose activation rows do not fit shared memory): units per tile and window, window count
  int fold;         // row folding (b200_qgemv_tiled_args::fold): N, K, batch above are the FOLDED view; outputs are unfolded
  int64_t Nout;     // true output features (fold * N)
  int Kin;          // true input features (K / fold)
  unsigned long long* trace;   // development: [cta][8] globaltimer stamps of this launch (NULL = off)
};

struct V4Plan {
  uint32_t bars, red, ssq, rstd, xsum, xs, tacc, ringA, lut, ringC;
  int nA, nC, S;
};

// Shared-memory carve-up by ABSOLUTE shared-window address (the table must sit at a 64 KB aligned address so that
// `byte << 8 | lane << 2 | table` is one PRMT).  Shared by the kernel and the host-side feasibility check.
__host__ __device__ inline V4Plan v4_plan(uint32_t b0, int stage, int red_bytes, bool int4, int B, int nkb, int xstride, int gap, int smem = kV4DynSmem, int extra = 0) {
  V4Plan pl{};
  uint32_t off = (b0 + 127u) & ~127u;
  pl.bars = off; off += 256;
  pl.red = off; off += (uint32_t)red_bytes;
  pl.ssq = off; off += kV4CW * 16 * 4;
  pl.rstd = off; off += 64;
  pl.xsum = off; if (int4) off += (((uint32_t)(B * nkb) + 3u) & ~3u) * 4;
  pl.xs = off; if (!gap) off += (((uint32_t)B * xstride * 2) + 127u) & ~127u;
  pl.tacc = off; off += (uint32_t)extra;
  off = (off + 127u) & ~127u;
  pl.ringA = off;
  const uint32_t end = b0 + (uint32_t)smem;
  if (int4) {
    pl.lut = 0; pl.ringC = end;
    pl.nA = off < end ? (int)((end - off) / stage) : 0; pl.nC = 0;
  } else {
    int best = -1;
    for (uint32_t cand = 0x10000u; cand + 0x10000u <= end; cand += 0x10000u) {
      if (cand < off) continue;
      const int nA = (int)((cand - off) / stage), nC = (int)((end - cand - 0x10000u) / stage);
      if (nA + nC > best) { best = nA + nC; pl.lut = cand; pl.ringC = cand + 0x10000u; pl.nA = nA; pl.nC = nC; }
    }
    if (best < 0) { pl.lut = 0; pl.nA = pl.nC = 0; }
  }
  pl.S = pl.nA + pl.nC < kV4MaxStages ? pl.nA + pl.nC : kV4MaxStages;
  return pl;
}

__device__ __forceinline__ void v4_stamp(const V4Params& p, int ev) {
  if (p.trace) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    p.trace[(size_t)blockIdx.x * 8 + ev] = t;
  }
}

__device__ __forceinline__ void v4_stamp_cyc(const V4Params& p, int ev, long long c0) {
  if (p.trace) p.trace[(size_t)blockIdx.x * 8 + ev] = (unsigned long long)(clock64() - c0);
}

__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float ld_cg_f32(const float* p) {
  float v;
  asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}

// MODE 0: batch 1 (one accumulator column), MODE 1: batch <= 8, MODE 2: batch <= 16
// XREG (batch 1, units per tile == consumer warps): a warp meets the same k-range in every unit, so its activation
// fragments (and int4 block sums) stay in registers for the whole launch instead of being re-read from shared memory
template <int RU, bool INT4, int MODE, bool XREG, int OCC = 1, bool WIN = false, bool ATT = false>
__global__ void __launch_bounds__(kV4Threads, OCC) qgemv_v4_kernel(const V4Params p) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  constexpr int NB = MODE == 2 ? 2 : 1;
  constexpr int COLS = 8 * NB;
  constexpr int REC = INT4 ? 1088 : 576;
  constexpr int BPR = INT4 ? 2 : 1;            // 64-wide k-blocks per record
  constexpr int UB = RU * REC;                 // unit bytes
  constexpr int STAGE = kV4CW * UB;
  constexpr int REDW = MODE == 0 ? 16 : 16 * COLS;   // floats per (buffer, warp)
  const int NBUF = MODE == 0 ? 4 : p.nbuf;
  // nf4 / fp4 at batch > 1: the reduction buffers live in the unused upper halves of the 256-byte table rows (batch 1 keeps its
  // activation row there), which frees 16 - 32 KB for the activation rows and the ring
  constexpr bool GAPRED = !INT4 && MODE != 0;
  constexpr int NV = MODE == 0 ? 1 : (16 * COLS) / 32;  // outputs per reducer lane and tile

  const uint32_t b0 = smem_u32(smem_raw);
  uint8_t* const sb = smem_raw - b0;           // generic pointer to shared-window address 0
  const int B = p.batch;
  // WIN: p.nkb / p.xstride describe ONE k-window; the per-tile accumulators of the reducer follow the activation rows
  const int tacc_bytes = WIN ? ((((int)((p.U / (p.ts * p.upt)) + (int)gridDim.x - 1) / (int)gridDim.x) * p.ts) * 16 * COLS * 4) : 0;
  const V4Plan pl = v4_plan(b0, STAGE, GAPRED ? 0 : NBUF * kV4CW * REDW * 4, INT4, B, p.nkb, p.xstride, p.gap, OCC == 2 ? kV4DynSmem2 : kV4DynSmem, tacc_bytes);
  const int S = pl.S;
  if (S < 2) __trap();                         // the host sized the launch so that this cannot happen
  const uint32_t full_a = pl.bars, empty_a = pl.bars + 64, rfull_a = pl.bars + 128, rempty_a = pl.bars + 160;
  auto stage_addr = [&](int s) -> uint32_t { return s < pl.nA ? pl.ringA + s * STAGE : pl.ringC + (s - pl.nA) * STAGE; };

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int G = gridDim.x, cta = blockIdx.x;
  const int upt = p.upt, TS = p.ts;
  // first unit of CTA c's range: equal unit counts (split) or equal entity counts (whole tiles / gate-up pairs)
  auto range_start = [&](int c) -> int {
    if (p.split) return (int)(((int64_t)p.U * c) / G);
    const int upe = TS * upt;
    return (int)(((int64_t)(p.U / upe) * c) / G) * upe;
  };
  const int u0 = range_start(cta), u1 = range_start(cta + 1);
  const int t_first = u0 / upt, t_last = (u1 - 1) / upt;   // flat tile range of this CTA (u1 > u0: the host keeps G <= entities / units)
  // this consumer warp's first unit (index arithmetic with divisions: done here, off the post-dependency critical path)
  const int u_first = u0 + warp, tf_first = u_first / upt, ku_first = u_first - tf_first * upt;

  if (tid == 0) {
    v4_stamp(p, 0);
    for (int s = 0; s < kV4MaxStages; ++s) { mbar_init(full_a + s * 8, 1); mbar_init(empty_a + s * 8, kV4CW); }
    for (int s = 0; s < 4; ++s) { mbar_init(rfull_a + s * 8, kV4CW); mbar_init(rempty_a + s * 8, 1); }
    mbar_init(pl.bars + 192, 1);   // activation rows copied by cp.async.bulk (batch > 1 without norm / delta)
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  if (warp == kV4CW) {
    // ================================ PRODUCER WARP (one elected lane) ==========================================
    if (lane == 0) {
      const uint64_t pol = make_evict_first_policy();
      int n = 0, ps = 0;
      uint32_t pph = 1;           // parity to wait for on the slot's EMPTY barrier (first lap: no wait)
      if constexpr (WIN) {
        // window-major order: for every k-window, the window's units of every tile of this CTA; a stage never straddles
        // a window boundary (the consumers swap the activation rows there)
        const int ntl = t_last - t_first + 1;
        for (int w = 0; w < p.nwin; ++w) {
          const int wl = min(p.wu, upt - w * p.wu), tot = ntl * wl;
          for (int j0 = 0; j0 < tot; j0 += kV4CW, ++n) {
            if (n >= S) mbar_wait(empty_a + ps * 8, pph);
            const int cnt = min(kV4CW, tot - j0);
            const uint32_t bar = full_a + ps * 8;
            uint32_t dst = stage_addr(ps);
            mbar_expect_tx(bar, (uint32_t)cnt * UB);
            int j = j0, rem = cnt;
            while (rem > 0) {
              const int tl = j / wl, kuw = j - tl * wl;
              const int run = min(rem, wl - kuw);
              const int tf = t_first + tl;
              const uint8_t* src = (TS == 2 && (tf & 1)) ? p.tB : p.tA;
              const int mt = TS == 2 ? (tf >> 1) : tf;
              bulk_g2s(dst, src + ((int64_t)mt * upt + w * p.wu + kuw) * UB, (uint32_t)run * UB, bar, pol);
              dst += (uint32_t)run * UB;
              j += run; rem -= run;
            }
            if (++ps == S) { ps = 0; pph ^= 1u; }
          }
        }
      } else
      for (int u = u0; u < u1; ++n) {
        if (n >= S) mbar_wait(empty_a + ps * 8, pph);
        const int cnt = min(kV4CW, u1 - u);
        const uint32_t bar = full_a + ps * 8;
        uint32_t dst = stage_addr(ps);
        mbar_expect_tx(bar, (uint32_t)cnt * UB);
        int v = u, rem = cnt;
        while (rem > 0) {          // contiguous pieces: a piece ends where the flat tile (and, for SwiGLU, the matrix) changes
          const int tf = v / upt, ku = v - tf * upt;
          const int run = min(rem, upt - ku);
          const uint8_t* src = (TS == 2 && (tf & 1)) ? p.tB : p.tA;
          const int mt = TS == 2 ? (tf >> 1) : tf;
          bulk_g2s(dst, src + ((int64_t)mt * upt + ku) * UB, (uint32_t)run * UB, bar, pol);
          dst += (uint32_t)run * UB;
          v += run; rem -= run;
        }
        u += cnt;
        if (++ps == S) { ps = 0; pph ^= 1u; }
      }
      if (p.debug != 3 && !ATT) v4_stamp(p, 6);
      if (p.kvh_k) {
        // this launch's weight stream is issued: ask the L2 for the filled rows of the K / V cache the attention kernel reads
        // next (one plane per CTA).  pos may lag by one step here (no dependency wait on this path): it is only a hint
        for (int c = cta; c < 2 * p.kvh_chunks; c += G) {
          long long pos = p.kvh_i64 ? *reinterpret_cast<const volatile long long*>(p.kvh_pos) : (long long)*reinterpret_cast<const volatile int*>(p.kvh_pos);
          if (pos > p.kvh_max) pos = p.kvh_max;
          const long long bytes = pos * p.kvh_row;
          const uint8_t* base = (c < p.kvh_chunks ? p.kvh_k : p.kvh_v) + (long long)(c % p.kvh_chunks) * p.kvh_stride;
          for (long long o = 0; o < bytes; o += 65536) prefetch_l2_bulk(base + o, (uint32_t)(bytes - o < 65536 ? ((bytes - o + 15) & ~15LL) : 65536));
        }
      }
    }
    griddep_launch();
    return;
  }

  float* red = reinterpret_cast<float*>(sb + pl.red);
  // float L of the reduction buffers (buffer-major, then warp, then output): plain array, or 32 floats per table-row gap
  auto red_at = [&](int L) -> float* {
    if constexpr (GAPRED) return reinterpret_cast<float*>(sb + pl.lut + 128u + (uint32_t)(L >> 5) * 256u + (uint32_t)(L & 31) * 4u);
    else return red + L;
  };

  if (warp == kV4CW + 1) {
    // ================================ REDUCER WARP ================================================================
    griddep_launch();
    griddep_wait();
    uint32_t* wsflag = reinterpret_cast<uint32_t*>(p.ws + 64);
    const uint32_t stag = p.tp_send_seq ? (__ldcg(p.tp_send_seq) % 65535u) + 1u : 0u;
    // STORE / RESIDUAL epilogues for one (row, batch column): `addend` = bias (+ nothing) and `res` = the residual value, both
    // loaded by the caller for ALL of a lane's outputs before the first store (y may alias residual, so the compiler cannot
    // hoist them itself: eight dependent L2 round trips per tile made the reducer the longest warp of a batched launch)
    auto out_index = [&](int64_t& n, int& c) {
      if (p.fold > 1) {  // folded view: row n holds the true rows n * fold + h, column c = batch row * fold + h
        n = n * p.fold + (c % p.fold);
        c /= p.fold;
      }
      return (int64_t)c * p.Nout + n;
    };
    auto store_at = [&](int64_t at, float v, float addend, float res) {
      float h = v + addend;
      if (p.epilogue == B200_EPI_RESIDUAL) h = round_bf16(h) + res;
      const __nv_bfloat16 hb = __float2bfloat16_rn(h);
      if (p.y) p.y[at] = hb;
      if (p.tp_send_seq) {  // row-parallel partial: one self-validating packet straight into every rank's slot over NVLink
        const uint32_t pk = (stag << 16) | (uint32_t)bf16_bits(h);
        for (int r = 0; r < p.tp_world; ++r)
          asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p.tp_peer_y[r] + at), "r"(pk) : "memory");
      }
    };
    // epilogue of a complete entity (tile, or gate tile + up tile): bias / residual / SwiGLU / tensor-parallel packets
    auto finish = [&](int e, float (&val)[2][NV]) {
    if (p.epilogue == B200_EPI_SWIGLU) {
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int o = MODE == 0 ? lane : 4 * lane + (i & 3) + 128 * (i >> 2);   // batch > 1: lane owns four consecutive outputs per 128
        const int row = MODE == 0 ? (lane & 15) : o / COLS, c = MODE == 0 ? 0 : o % COLS;
        const int64_t nrow = (int64_t)e * 16 + row;
        if ((MODE != 0 || lane < 16) && c < B && nrow < p.N) {
          const float a = round_bf16(val[0][i] + (p.bias ? __bfloat162float(p.bias[nrow]) : 0.f));
          const float b = round_bf16(val[1][i] + (p.bias2 ? __bfloat162float(p.bias2[nrow]) : 0.f));
          p.y[(int64_t)c * p.N + nrow] = __float2bfloat16_rn(silu_bf16_s(a) * b);
        }
      }
    } else {
      int64_t at[NV];
      float addend[NV], res[NV];
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int o = MODE == 0 ? lane : 4 * lane + (i & 3) + 128 * (i >> 2);   // batch > 1: lane owns four consecutive outputs per 128
        const int row = MODE == 0 ? (lane & 15) : o / COLS;
        int c = MODE == 0 ? 0 : o % COLS;
        int64_t nrow = (int64_t)e * 16 + row;
        const bool ok = (MODE != 0 || lane < 16) && c < B && nrow < p.N;
        at[i] = ok ? out_index(nrow, c) : -1;
        addend[i] = ok && p.bias ? __bfloat162float(p.bias[nrow]) : 0.f;
      }
#pragma unroll
      for (int i = 0; i < NV; ++i)
        res[i] = at[i] >= 0 && p.epilogue == B200_EPI_RESIDUAL ? __bfloat162float(p.residual[at[i]]) : 0.f;
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        if (at[i] >= 0) store_at(at[i], val[0][i], addend[i], res[i]);
      }
    }
      };
    const int e_first = t_first / TS, e_last = t_last / TS;
    if constexpr (WIN) {
      // k-windows: the partial sums of (window, tile) arrive in window-major order and meet in this warp's per-tile
      // accumulators (fixed order: deterministic); an entity is finished when its last window has arrived
      const int ntl = t_last - t_first + 1, nq = p.nwin * ntl;
      float* tacc = reinterpret_cast<float*>(sb + pl.tacc);
      for (int q = 0; q < nq; ++q) {
        const int w = q / ntl, tl = q - w * ntl;
        const int buf = q % NBUF;
        mbar_wait(rfull_a + buf * 8, (uint32_t)(q / NBUF) & 1u);
        const int rb = buf * (kV4CW * REDW);
        float val[2][NV];
#pragma unroll
        for (int h = 0; h < NV / 4; ++h) {
          float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int ww = 0; ww < kV4CW; ++ww) {
            const float4 q4 = *reinterpret_cast<const float4*>(red_at(rb + ww * REDW + 128 * h + 4 * lane));
            v.x += q4.x; v.y += q4.y; v.z += q4.z; v.w += q4.w;
          }
          val[1][4 * h] = v.x; val[1][4 * h + 1] = v.y; val[1][4 * h + 2] = v.z; val[1][4 * h + 3] = v.w;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(rempty_a + buf * 8);
        float* ta = tacc + tl * (16 * COLS);
        if (w > 0) {
#pragma unroll
          for (int i = 0; i < NV; ++i) val[1][i] += ta[lane + 32 * i];
        }
        const int tf = t_first + tl;
        const bool last = w == p.nwin - 1;
        if (!last || (TS == 2 && !(tf & 1))) {   // not complete yet (or the gate tile waiting for its up tile): park it
#pragma unroll
          for (int i = 0; i < NV; ++i) ta[lane + 32 * i] = val[1][i];
          continue;
        }
        if (TS == 2) {
#pragma unroll
          for (int i = 0; i < NV; ++i) val[0][i] = (ta - 16 * COLS)[lane + 32 * i];
        } else {
#pragma unroll
          for (int i = 0; i < NV; ++i) val[0][i] = val[1][i];
        }
        finish(tf / TS, val);
      }
      if (lane == 0) v4_stamp(p, 5);
      return;
    }
    if (lane == 0 && p.debug != 3 && !ATT) v4_stamp(p, 7);
    int j = 0;  // tiles reduced so far (reduction-buffer sequence)
    for (int e = e_first; e <= e_last; ++e) {
      float val[2][NV];
#pragma unroll
      for (int w = 0; w < 2; ++w) {
#pragma unroll
        for (int i = 0; i < NV; ++i) val[w][i] = 0.f;
      }
#pragma unroll
      for (int which = 0; which < 2; ++which) {
        const int tf = e * TS + which;
        if (which >= TS || tf < t_first || tf > t_last) continue;
        const int buf = j % NBUF;
        mbar_wait(rfull_a + buf * 8, (uint32_t)(j / NBUF) & 1u);
        const int rb = buf * (kV4CW * REDW);
        if constexpr (MODE == 0) {
          const int row = lane & 15, half = lane >> 4;
          float v = 0.f;
#pragma unroll
          for (int w = 0; w < kV4CW / 2; ++w) v += red[rb + (half * (kV4CW / 2) + w) * REDW + row];
          v += __shfl_xor_sync(0xffffffffu, v, 16);   // (warps 0..7) + (warps 8..15): same value in both halves
          val[which][0] = v;
        } else {
#pragma unroll
          for (int h = 0; h < NV / 4; ++h) {   // 128-bit loads: 2 x 16 (batch <= 8: 1 x 16) instead of 128 scalar ones
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int w = 0; w < kV4CW; ++w) {
              const float4 q4 = *reinterpret_cast<const float4*>(red_at(rb + w * REDW + 128 * h + 4 * lane));
              v.x += q4.x; v.y += q4.y; v.z += q4.z; v.w += q4.w;
            }
            val[which][4 * h] = v.x; val[which][4 * h + 1] = v.y; val[which][4 * h + 2] = v.z; val[which][4 * h + 3] = v.w;
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(rempty_a + buf * 8);
        ++j;
      }
      const int ent_u0 = e * TS * upt, ent_u1 = ent_u0 + TS * upt;
      if (ent_u0 < u0) {
        // CONTRIBUTOR: the entity's first unit belongs to a lower CTA; publish this CTA's partial sums
        float* wd = reinterpret_cast<float*>(p.ws + kV4WsData + (size_t)cta * kV4WsPerCta);
#pragma unroll
        for (int w = 0; w < 2; ++w) {
          if (w >= TS) continue;
          if constexpr (MODE == 0) {
            if (lane < 16) wd[w * 16 + lane] = val[w][0];
          } else {
#pragma unroll
            for (int i = 0; i < NV; ++i) wd[w * (16 * COLS) + lane + 32 * i] = val[w][i];
          }
        }
        __threadfence();
        __syncwarp();
        if (lane == 0) asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(wsflag + cta), "r"(1u) : "memory");
        continue;
      }
      if (ent_u1 > u1) {
        // OWNER of an entity cut by the range boundary: add the partials of the CTAs holding the rest, in CTA order
        for (int c = cta + 1; c < G && range_start(c) < ent_u1; ++c) {
          if (lane == 0) {
            const long long t0 = clock64();
            while (ld_acquire_gpu(wsflag + c) == 0u) {
              if (clock64() - t0 > 4000000000LL) { *reinterpret_cast<volatile uint32_t*>(p.ws) = 1u; break; }  // error word, no hang
            }
          }
          __syncwarp();
          const float* wd = reinterpret_cast<const float*>(p.ws + kV4WsData + (size_t)c * kV4WsPerCta);
#pragma unroll
          for (int w = 0; w < 2; ++w) {
            if (w >= TS) continue;
            if constexpr (MODE == 0) {
              val[w][0] += ld_cg_f32(wd + w * 16 + (lane & 15));
            } else {
#pragma unroll
              for (int i = 0; i < NV; ++i) val[w][i] += ld_cg_f32(wd + w * (16 * COLS) + lane + 32 * i);
            }
          }
          __syncwarp();
          if (lane == 0) *reinterpret_cast<volatile uint32_t*>(wsflag + c) = 0u;  // self-cleaning: ready for the next launch
        }
      }
      finish(e, val);
    }
    if constexpr (ATT) {
      // this CTA's row tiles of q / k / v are in memory: tell the attention sub-blocks, ONE fence for all of them (a fence
      // per tile cost ~0.6 us each); the tiles of a CTA span at most a few heads
      __threadfence();
      __syncwarp();
      if (lane == 0) {
        int e = e_first;
        while (e <= e_last) {
          const int h = e / p.att_tph, e_end = min(e_last + 1, (h + 1) * p.att_tph);
          asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p.att_cnt + h), "r"((uint32_t)(e_end - e)) : "memory");
          e = e_end;
        }
      }
    }
    if (lane == 0) v4_stamp(p, 5);
    return;
  }

  // ================================== CONSUMER WARPS ===============================================================
  const int g = lane >> 2, t = lane & 3;
  float* s_ssq = reinterpret_cast<float*>(sb + pl.ssq);
  float* s_rstd = reinterpret_cast<float*>(sb + pl.rstd);
  float* xsum_s = reinterpret_cast<float*>(sb + pl.xsum);
  const uint32_t lut_a = pl.lut;
  if (!INT4) {
    // table entry e = warp + 16 k, lane slot `lane` at byte e * 256 + lane * 4: element 0 (low half) = code[e >> 4] = code[k],
    // element 1 = code[e & 15] = code[warp] (even element = HIGH nibble).  Sixteen independent constant loads + stores per thread
    uint32_t* lut = reinterpret_cast<uint32_t*>(sb + lut_a);
    const float* cb = p.qtype == B200_Q_NF4 ? kNF4 : kFP4;
    const float mine = cb[warp];
#pragma unroll
    for (int k = 0; k < 16; ++k) lut[(warp + 16 * k) * 64 + lane] = pack_bf16(cb[k], mine);
  }
  // 16-byte chunk c of activation row b in shared memory (batch 1: the row lives in the upper halves of the table rows)
  auto xs_chunk = [&](int b, int c) -> uint32_t {
    return p.gap ? lut_a + (uint32_t)(c >> 3) * 256u + 128u + (uint32_t)(c & 7) * 16u : pl.xs + (uint32_t)(b * p.xstride) * 2u + (uint32_t)c * 16u;
  };
  // RMSNorm weights are constants: fetch this thread's first chunks BEFORE waiting for the previous kernel
  const int chunks = (int)(p.K >> 3);
  uint4 nwreg[3] = {make_uint4(0u, 0u, 0u, 0u), make_uint4(0u, 0u, 0u, 0u), make_uint4(0u, 0u, 0u, 0u)};
  if (p.norm_w) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const int c = tid + i * kV4Consumers;
      if (c < chunks) nwreg[i] = *reinterpret_cast<const uint4*>(p.norm_w + c * 8);
    }
  }
  griddep_launch();
  griddep_wait();
  if (tid == 0) v4_stamp(p, 1);
  const long long c_rel = clock64();

  // ---- tensor-parallel consumer: the branch output is the sum of every rank's partial, written into this rank's slots by
  // the peers' epilogues (and its own) as self-validating packets; the expected tag comes from the buffer's local
  // sequence counter (last bumped at least two launches ago)
  uint32_t rtag = 0;
  if (p.tp_slots) rtag = (__ldcg(p.tp_recv_seq) % 65535u) + 1u;
  // ---- activation prologue: (x + delta) -> optional RMSNorm -> bf16 rows in shared memory -------------------------
  const bool designated = cta == 0;
  constexpr int UK = RU * BPR * 64;            // k elements per unit
  // WIN: stage window w of the activation rows: (x + delta) -> optional RMSNorm with the row's rstd -> bf16 in shared memory
  auto add_bf16x8 = [](uint4 v, const uint4 dv) -> uint4 {
    v.x = pack_bf16(bf16lo(v.x) + bf16lo(dv.x), bf16hi(v.x) + bf16hi(dv.x));
    v.y = pack_bf16(bf16lo(v.y) + bf16lo(dv.y), bf16hi(v.y) + bf16hi(dv.y));
    v.z = pack_bf16(bf16lo(v.z) + bf16lo(dv.z), bf16hi(v.z) + bf16hi(dv.z));
    v.w = pack_bf16(bf16lo(v.w) + bf16lo(dv.w), bf16hi(v.w) + bf16hi(dv.w));
    return v;
  };
  auto stage_window = [&](int w) {
    const int k0 = w * p.wu * UK;
    const int kw = min(p.wu * UK, (int)p.K - k0), wch = kw >> 3;
    const int total = B * wch;                 // 16-byte chunks of the window, all rows: four loads in flight per thread
    if (!p.norm_w && !p.delta) {
      // plain rows (the 16-token prompt prefill, row-parallel Linears): one bulk copy per row, no per-element instructions
      if (tid == 0) {
        mbar_expect_tx(pl.bars + 192, (uint32_t)(B * kw * 2));
        for (int b = 0; b < B; ++b)
          bulk_g2s_plain(pl.xs + (uint32_t)(b * p.xstride) * 2u, p.x + (int64_t)b * p.K + k0, (uint32_t)(kw * 2), pl.bars + 192);
      }
      mbar_wait(pl.bars + 192, (uint32_t)w & 1u);
    } else {
    for (int i0 = tid; i0 < total; i0 += 4 * kV4Consumers) {
      uint4 v[4], dv[4], nw[4];
      int bb[4], cc[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * kV4Consumers;
        bb[u] = i / wch; cc[u] = i - bb[u] * wch;
        if (i < total) {
          v[u] = *reinterpret_cast<const uint4*>(p.x + (int64_t)bb[u] * p.K + k0 + cc[u] * 8);
          if (p.delta) dv[u] = *reinterpret_cast<const uint4*>(p.delta + (int64_t)bb[u] * p.K + k0 + cc[u] * 8);
          if (p.norm_w) nw[u] = *reinterpret_cast<const uint4*>(p.norm_w + k0 + cc[u] * 8);
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (i0 + u * kV4Consumers < total) {
          uint4 x4 = v[u];
          if (p.delta) x4 = add_bf16x8(x4, dv[u]);
          if (p.norm_w) {
            const float rstd = s_rstd[bb[u]];
            const uint4 n4 = nw[u];
            x4.x = pack_bf16(bf16lo(n4.x) * (bf16lo(x4.x) * rstd), bf16hi(n4.x) * (bf16hi(x4.x) * rstd));
            x4.y = pack_bf16(bf16lo(n4.y) * (bf16lo(x4.y) * rstd), bf16hi(n4.y) * (bf16hi(x4.y) * rstd));
            x4.z = pack_bf16(bf16lo(n4.z) * (bf16lo(x4.z) * rstd), bf16hi(n4.z) * (bf16hi(x4.z) * rstd));
            x4.w = pack_bf16(bf16lo(n4.w) * (bf16lo(x4.w) * rstd), bf16hi(n4.w) * (bf16hi(x4.w) * rstd));
          }
          *reinterpret_cast<uint4*>(sb + pl.xs + (uint32_t)(bb[u] * p.xstride) * 2u + (uint32_t)cc[u] * 16u) = x4;
        }
      }
    }
    asm volatile("bar.sync 1, %0;" ::"n"(kV4Consumers) : "memory");
    }
    if (INT4) {   // per (row, k-block of the window) activation sums for the zero-point term (blocks past a short last window: unused)
      xsum_blocks<kV4Consumers>(reinterpret_cast<const __nv_bfloat16*>(sb + pl.xs), p.xstride, B, p.nkb, xsum_s, tid);
      asm volatile("bar.sync 1, %0;" ::"n"(kV4Consumers) : "memory");
    }
  };
  if constexpr (WIN) {
    // pass 1 over the full rows: publish x + delta once, sum of squares -> rstd per row; the rows themselves are staged
    // window by window in the main loop (same arithmetic and rounding points as the one-window path)
    if (p.norm_w || (p.delta && p.x_out)) {
      for (int b = 0; b < B; ++b) {
        float ss = 0.f;
        for (int c0 = tid; c0 < chunks; c0 += 4 * kV4Consumers) {   // four loads in flight per thread
          uint4 v[4], dv[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int c = c0 + u * kV4Consumers;
            if (c < chunks) {
              v[u] = *reinterpret_cast<const uint4*>(p.x + (int64_t)b * p.K + c * 8);
              if (p.delta) dv[u] = *reinterpret_cast<const uint4*>(p.delta + (int64_t)b * p.K + c * 8);
            }
          }
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int c = c0 + u * kV4Consumers;
            if (c < chunks) {
              uint4 x4 = v[u];
              if (p.delta) {
                x4 = add_bf16x8(x4, dv[u]);
                if (p.x_out && designated) *reinterpret_cast<uint4*>(p.x_out + (int64_t)b * p.K + c * 8) = x4;
              }
              const uint32_t w4[4] = {x4.x, x4.y, x4.z, x4.w};
#pragma unroll
              for (int i = 0; i < 4; ++i) { const float a2 = bf16lo(w4[i]), c2 = bf16hi(w4[i]); ss = fmaf(a2, a2, ss); ss = fmaf(c2, c2, ss); }
            }
          }
        }
        if (p.norm_w) {
          ss = warp_sum(ss);
          if (lane == 0) s_ssq[warp * 16 + b] = ss;
        }
      }
      if (p.norm_w) {
        asm volatile("bar.sync 1, %0;" ::"n"(kV4Consumers) : "memory");
        if (tid < B) {
          float tot = 0.f;
#pragma unroll
          for (int w = 0; w < kV4CW; ++w) tot += s_ssq[w * 16 + tid];
          s_rstd[tid] = rsqrtf(tot / (float)p.K + p.eps);
        }
      }
    }
  } else
  if (B == 1 && !p.tp_slots && chunks <= 3 * kV4Consumers) {
    // batch-1 fast path (the decode chain): every load of the row is in flight at once, ONE pass over registers, the
    // sum of squares meets in shared memory with a single barrier and every thread derives rstd itself (same
    // summation order and arithmetic as the general path below, so the two are bit-identical)
    uint4 xv[3], dv[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const int c = tid + i * kV4Consumers;
      if (c < chunks) xv[i] = *reinterpret_cast<const uint4*>(p.x + c * 8);
    }
    if (p.delta) {
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        const int c = tid + i * kV4Consumers;
        if (c < chunks) dv[i] = *reinterpret_cast<const uint4*>(p.delta + c * 8);
      }
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        const int c = tid + i * kV4Consumers;
        if (c < chunks) {
          uint4 v = xv[i];
          v.x = pack_bf16(bf16lo(v.x) + bf16lo(dv[i].x), bf16hi(v.x) + bf16hi(dv[i].x));
          v.y = pack_bf16(bf16lo(v.y) + bf16lo(dv[i].y), bf16hi(v.y) + bf16hi(dv[i].y));
          v.z = pack_bf16(bf16lo(v.z) + bf16lo(dv[i].z), bf16hi(v.z) + bf16hi(dv[i].z));
          v.w = pack_bf16(bf16lo(v.w) + bf16lo(dv[i].w), bf16hi(v.w) + bf16hi(dv[i].w));
          if (p.x_out && designated) *reinterpret_cast<uint4*>(p.x_out + c * 8) = v;
          xv[i] = v;
        }
      }
    }
    if (p.norm_w) {
      float ss = 0.f;
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        if (tid + i * kV4Consumers < chunks) {
          const uint32_t w4[4] = {xv[i].x, xv[i].y, xv[i].z, xv[i].w};
#pragma unroll
          for (int q = 0; q < 4; ++q) { const float a = bf16lo(w4[q]), c2 = bf16hi(w4[q]); ss = fmaf(a, a, ss); ss = fmaf(c2, c2, ss); }
        }
      }
      ss = warp_sum(ss);
      if (lane == 0) s_ssq[warp] = ss;
      asm volatile("bar.sync 1, %0;" ::"n"(kV4Consumers) : "memory");
      const float4 q0 = *reinterpret_cast<const float4*>(s_ssq), q1 = *reinterpret_cast<const float4*>(s_ssq + 4),
                   q2 = *reinterpret_cast<const float4*>(s_ssq + 8), q3 = *reinterpret_cast<const float4*>(s_ssq + 12);
      float tot = 0.f;
      tot += q0.x; tot += q0.y; tot += q0.z; tot += q0.w; tot += q1.x; tot += q1.y; tot += q1.z; tot += q1.w;
      tot += q2.x; tot += q2.y; tot += q2.z; tot += q2.w; tot += q3.x; tot += q3.y; tot += q3.z; tot += q3.w;
      const float rstd = rsqrtf(tot / (float)p.K + p.eps);
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        if (tid + i * kV4Consumers < chunks) {
          const uint4 nw = nwreg[i];
          uint4 v = xv[i];
          // weight * (x * rstd), fp32, one cast (rmsnorm.py:24-25)
          v.x = pack_bf16(bf16lo(nw.x) * (bf16lo(v.x) * rstd), bf16hi(nw.x) * (bf16hi(v.x) * rstd));
          v.y = pack_bf16(bf16lo(nw.y) * (bf16lo(v.y) * rstd), bf16hi(nw.y) * (bf16hi(v.y) * rstd));
          v.z = pack_bf16(bf16lo(nw.z) * (bf16lo(v.z) * rstd), bf16hi(nw.z) * (bf16hi(v.z) * rstd));
          v.w = pack_bf16(bf16lo(nw.w) * (bf16lo(v.w) * rstd), bf16hi(nw.w) * (bf16hi(v.w) * rstd));
          xv[i] = v;
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const int c = tid + i * kV4Consumers;
      if (c < chunks) *reinterpret_cast<uint4*>(sb + xs_chunk(0, c)) = xv[i];
    }
    if (tid == 0 && p.debug == 3) v4_stamp_cyc(p, 7, c_rel);
  } else {
    if (!p.tp_slots && p.fold <= 1 && !p.norm_w && !p.delta && !p.gap) {
      // batch > 1, plain rows: one bulk copy per row straight into shared memory (no per-element instructions)
      if (tid == 0) {
        mbar_expect_tx(pl.bars + 192, (uint32_t)(B * (int)p.K * 2));
        for (int b = 0; b < B; ++b)
          bulk_g2s_plain(pl.xs + (uint32_t)(b * p.xstride) * 2u, p.x + (int64_t)b * p.K, (uint32_t)((int)p.K * 2), pl.bars + 192);
      }
      mbar_wait(pl.bars + 192, 0u);
    } else if (!p.tp_slots && p.fold <= 1) {
      // batch > 1: the (row, chunk) items of this thread in row-major order, four global loads in flight at a time (a plain
      // row loop serialises one L2 round trip per row: the shared-memory stores in between keep ptxas from hoisting loads)
      const int ncm = (chunks + kV4Consumers - 1) / kV4Consumers, items = B * ncm;
      float ss = 0.f;
      for (int t0 = 0; t0 < items; t0 += 4) {
        uint4 v[4], dv[4];
        int bb[4], cc[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int t = t0 + u;
          bb[u] = t / ncm; cc[u] = tid + (t - bb[u] * ncm) * kV4Consumers;
          if (t < items && cc[u] < chunks) {
            v[u] = *reinterpret_cast<const uint4*>(p.x + (int64_t)bb[u] * p.K + cc[u] * 8);
            if (p.delta) dv[u] = *reinterpret_cast<const uint4*>(p.delta + (int64_t)bb[u] * p.K + cc[u] * 8);
          }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int t = t0 + u;
          if (t < items) {
            if (cc[u] < chunks) {
              uint4 x4 = v[u];
              if (p.delta) {
                x4 = add_bf16x8(x4, dv[u]);
                if (p.x_out && designated) *reinterpret_cast<uint4*>(p.x_out + (int64_t)bb[u] * p.K + cc[u] * 8) = x4;
              }
              *reinterpret_cast<uint4*>(sb + xs_chunk(bb[u], cc[u])) = x4;
              const uint32_t w4[4] = {x4.x, x4.y, x4.z, x4.w};
#pragma unroll
              for (int i = 0; i < 4; ++i) { const float a2 = bf16lo(w4[i]), c2 = bf16hi(w4[i]); ss = fmaf(a2, a2, ss); ss = fmaf(c2, c2, ss); }
            }
            if (t - bb[u] * ncm == ncm - 1) {   // last item of the row (uniform across the CTA): the row's sum of squares
              if (p.norm_w) {
                ss = warp_sum(ss);
                if (lane == 0) s_ssq[warp * 16 + bb[u]] = ss;
              }
              ss = 0.f;
            }
          }
        }
      }
    } else {
    const uint32_t* tp_src = p.tp_slots;
    int tp_nw = p.tp_world;
    if (p.tp_slots && p.tp_reduced) {
      // two-stage exchange, stage 1: this CTA sums ITS slice of the vector over the ranks (fp32, rank order, one rounding to
      // bf16: the arithmetic of the one-stage path) and republishes it locally as packets with the same tag; stage 2 below
      // then reads one buffer instead of `world` slots
      const int total = B * chunks;
      const int i0 = (int)(((int64_t)total * cta) / G), i1 = (int)(((int64_t)total * (cta + 1)) / G);
      for (int i = i0 + tid; i < i1; i += kV4Consumers) {
        float acc8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        for (int r = 0; r < p.tp_world; ++r) {
          const uint32_t* sp = p.tp_slots + r * p.tp_slot_stride + (int64_t)i * 8;
          uint4 pa, pb;
          const long long t0 = clock64();
          for (;;) {
            asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(pa.x), "=r"(pa.y), "=r"(pa.z), "=r"(pa.w) : "l"(sp) : "memory");
            asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(pb.x), "=r"(pb.y), "=r"(pb.z), "=r"(pb.w) : "l"(sp + 4) : "memory");
            const bool ok = (pa.x >> 16) == rtag && (pa.y >> 16) == rtag && (pa.z >> 16) == rtag && (pa.w >> 16) == rtag &&
                            (pb.x >> 16) == rtag && (pb.y >> 16) == rtag && (pb.z >> 16) == rtag && (pb.w >> 16) == rtag;
            if (ok) break;
            if (clock64() - t0 > 4000000000LL) { *reinterpret_cast<volatile uint32_t*>(p.ws) = 2u; break; }
          }
          acc8[0] += bf16lo(pa.x); acc8[1] += bf16lo(pa.y); acc8[2] += bf16lo(pa.z); acc8[3] += bf16lo(pa.w);
          acc8[4] += bf16lo(pb.x); acc8[5] += bf16lo(pb.y); acc8[6] += bf16lo(pb.z); acc8[7] += bf16lo(pb.w);
        }
        const uint32_t tg = rtag << 16;
        uint32_t* dp = p.tp_reduced + (int64_t)i * 8;
        asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(dp), "r"(tg | bf16_bits(acc8[0])), "r"(tg | bf16_bits(acc8[1])),
                     "r"(tg | bf16_bits(acc8[2])), "r"(tg | bf16_bits(acc8[3])) : "memory");
        asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(dp + 4), "r"(tg | bf16_bits(acc8[4])), "r"(tg | bf16_bits(acc8[5])),
                     "r"(tg | bf16_bits(acc8[6])), "r"(tg | bf16_bits(acc8[7])) : "memory");
      }
      tp_src = p.tp_reduced;
      tp_nw = 1;
    }
    for (int b = 0; b < B; ++b) {
      float ss = 0.f;
      for (int c = tid; c < chunks; c += kV4Consumers) {
        uint4 v;
        if (p.fold > 1) {  // folded row b = (true row b / fold, segment h): the activation in segment h, zeros elsewhere
          const int h = b % p.fold, e0 = c * 8 - h * p.Kin;
          v = (e0 >= 0 && e0 < p.Kin) ? *reinterpret_cast<const uint4*>(p.x + (int64_t)(b / p.fold) * p.Kin + e0) : make_uint4(0u, 0u, 0u, 0u);
        } else {
          v = *reinterpret_cast<const uint4*>(p.x + (int64_t)b * p.K + c * 8);
        }
        if (p.delta || p.tp_slots) {
          uint4 dv;
          if (p.tp_slots) {  // sum of the ranks' partials in rank order (fp32), rounded once to bf16 like an all-reduced tensor
            float acc8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            for (int r = 0; r < tp_nw; ++r) {
              const uint32_t* sp = tp_src + r * p.tp_slot_stride + (int64_t)b * p.K + c * 8;
              uint4 pa, pb;
              const long long t0 = clock64();
              for (;;) {   // poll until all eight packets carry this hand-off's tag
                asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(pa.x), "=r"(pa.y), "=r"(pa.z), "=r"(pa.w) : "l"(sp) : "memory");
                asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(pb.x), "=r"(pb.y), "=r"(pb.z), "=r"(pb.w) : "l"(sp + 4) : "memory");
                const bool ok = (pa.x >> 16) == rtag && (pa.y >> 16) == rtag && (pa.z >> 16) == rtag && (pa.w >> 16) == rtag &&
                                (pb.x >> 16) == rtag && (pb.y >> 16) == rtag && (pb.z >> 16) == rtag && (pb.w >> 16) == rtag;
                if (ok) break;
                if (clock64() - t0 > 4000000000LL) {  // a peer died: raise the error word and carry on with what is there (no hang, no trap)
                  *reinterpret_cast<volatile uint32_t*>(p.ws) = 2u;
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
        *reinterpret_cast<uint4*>(sb + xs_chunk(b, c)) = v;
        const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
  #pragma unroll
        for (int i = 0; i < 4; ++i) { const float a = bf16lo(w4[i]), c2 = bf16hi(w4[i]); ss = fmaf(a, a, ss); ss = fmaf(c2, c2, ss); }
      }
      if (p.norm_w) {
        ss = warp_sum(ss);
        if (lane == 0) s_ssq[warp * 16 + b] = ss;
      }
    }
    }
    if (tid == 0 && p.debug == 3) v4_stamp_cyc(p, 3, c_rel);
    if (p.norm_w) {
      asm volatile("bar.sync 1, %0;" ::"n"(kV4Consumers) : "memory");
      if (tid == 0 && p.debug == 3) v4_stamp_cyc(p, 6, c_rel);
      if (tid < B) {
        float tot = 0.f;
  #pragma unroll
        for (int w = 0; w < kV4CW; ++w) tot += s_ssq[w * 16 + tid];
        s_rstd[tid] = rsqrtf(tot / (float)p.K + p.eps);
      }
      asm volatile("bar.sync 1, %0;" ::"n"(kV4Consumers) : "memory");
      for (int b = 0; b < B; ++b) {
        const float rstd = s_rstd[b];
        int ci = 0;
        for (int c = tid; c < chunks; c += kV4Consumers, ++ci) {  // same thread owns the same chunk as above
          uint4* px = reinterpret_cast<uint4*>(sb + xs_chunk(b, c));
          uint4 v = *px;
          const uint4 nw = ci == 0 ? nwreg[0] : ci == 1 ? nwreg[1] : ci == 2 ? nwreg[2] : *reinterpret_cast<const uint4*>(p.norm_w + c * 8);
          // weight * (x * rstd), fp32, one cast (rmsnorm.py:24-25)
          v.x = pack_bf16(bf16lo(nw.x) * (bf16lo(v.x) * rstd), bf16hi(nw.x) * (bf16hi(v.x) * rstd));
          v.y = pack_bf16(bf16lo(nw.y) * (bf16lo(v.y) * rstd), bf16hi(nw.y) * (bf16hi(v.y) * rstd));
          v.z = pack_bf16(bf16lo(nw.z) * (bf16lo(v.z) * rstd), bf16hi(nw.z) * (bf16hi(v.z) * rstd));
          v.w = pack_bf16(bf16lo(nw.w) * (bf16lo(v.w) * rstd), bf16hi(nw.w) * (bf16hi(v.w) * rstd));
          *px = v;
        }
      }
      if (tid == 0 && p.debug == 3) v4_stamp_cyc(p, 7, c_rel);
    }
  }
  asm volatile("bar.sync 1, %0;" ::"n"(kV4Consumers) : "memory");  // activation rows and the table are published
  if (tid == 0) v4_stamp(p, 2);
  if (p.tp_slots && tid == 0) {  // last CTA to have read the slots bumps the buffer's sequence counter for its next use
    const uint32_t old = atomicAdd(p.tp_recv_done, 1u);
    if (old == (uint32_t)G - 1u) {
      *p.tp_recv_done = 0u;
      *p.tp_recv_seq = __ldcg(p.tp_recv_seq) + 1u;
    }
  }
  if (INT4 && !WIN) {       // per (batch row, k-block) activation sums for the zero-point term
    xsum_blocks<kV4Consumers>(reinterpret_cast<const __nv_bfloat16*>(sb + pl.xs), p.xstride, B, p.nkb, xsum_s, tid);
    asm volatile("bar.sync 1, %0;" ::"n"(kV4Consumers) : "memory");
  }

  // table address = table | byte << 8 | lane << 2, assembled by ONE PRMT (the table is 64 KB aligned): picks lane_full.b0,
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

  // activation fragment addresses: lane (g, t) feeds batch column g (+ 8) with k = 16 t .. 16 t + 15 of the k-block
  const int xrow0 = (MODE != 0 && g < B) ? g : 0;
  const int xrow1 = (g + 8 < B) ? g + 8 : 0;
  const uint32_t xkb = p.gap ? 256u : 128u;
  const uint32_t xa0 = (p.gap ? lut_a + 128u : pl.xs + (uint32_t)(xrow0 * p.xstride) * 2u) + (uint32_t)t * 32u;
  const uint32_t xa1 = pl.xs + (uint32_t)(xrow1 * p.xstride) * 2u + (uint32_t)t * 32u;

  // activation fragment of k-block kb: batch 1 reads 4 x 64 bit (the lanes of a column group read the same address: one
  // wavefront each; a 128-bit load costs four wavefronts even when it broadcasts), batch > 1 reads 2 x 128 bit
  auto load_x = [&](uint32_t (&xb)[NB][8], int kb) {
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
      const uint32_t xa = (nb == 0 ? xa0 : xa1) + (uint32_t)kb * xkb;
      if constexpr (MODE == 0) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
          asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(xb[nb][2 * i]), "=r"(xb[nb][2 * i + 1]) : "r"(xa + 8u * i));
      } else {
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(xb[nb][0]), "=r"(xb[nb][1]), "=r"(xb[nb][2]), "=r"(xb[nb][3]) : "r"(xa));
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(xb[nb][4]), "=r"(xb[nb][5]), "=r"(xb[nb][6]), "=r"(xb[nb][7]) : "r"(xa + 16u));
      }
    }
  };
  constexpr int XR = XREG ? RU * BPR : 1;
  uint32_t xreg[XR][NB][8];
  float xsreg[XR];
  if constexpr (XREG) {
    const int ku0 = ku_first;   // the k-unit of every unit this warp will meet (unit stride == units per tile)
#pragma unroll
    for (int i = 0; i < XR; ++i) {
      load_x(xreg[i], ku0 * XR + i);
      xsreg[i] = INT4 ? xsum_s[ku0 * XR + i] : 0.f;
    }
  }

  float acc[NB][4];
#pragma unroll
  for (int nb = 0; nb < NB; ++nb) {
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[nb][i] = 0.f;
  }

  // one 64-wide k-block of one tile: w = this lane's 16 B {row g bytes, row g + 8 bytes}; sc = scale words of rows g, g + 8;
  // kb = k-block in the row, xi = k-block within the unit
  auto compute = [&](const uint4 w, const uint2 sc, int kb, int xi) {
    uint32_t xl[NB][8];
    if constexpr (!XREG) load_x(xl, kb);
    uint32_t (&xb)[NB][8] = XREG ? xreg[XREG ? xi : 0] : xl;
    float pr[NB][4];
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
#pragma unroll
      for (int i = 0; i < 4; ++i) pr[nb][i] = 0.f;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t wa = (j < 2) ? w.x : w.y, wb = (j < 2) ? w.z : w.w;
      uint32_t a0, a1, a2, a3;
      if constexpr (INT4) {
        // the tiled words hold even codes in the low and odd codes in the high half-word: one shift + one LOP3 yields the
        // bf16 pair (128 + e_2p, 128 + e_2p+1); the +128 joins the zero-point term below
        const int sh = 8 * (j & 1);
        auto ex = [&](uint32_t x, int s) -> uint32_t {
          uint32_t r;
          asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(r) : "r"(s ? x >> s : x), "r"(i4mask), "r"(i4magic));
          return r;
        };
        a0 = ex(wa, sh); a1 = ex(wb, sh); a2 = ex(wa, sh + 4); a3 = ex(wb, sh + 4);
      } else {
        const uint32_t sel0 = (j & 1) ? 0x7624u : 0x7604u, sel1 = (j & 1) ? 0x7634u : 0x7614u;
        a0 = LUT(wa, sel0); a2 = LUT(wa, sel1); a1 = LUT(wb, sel0); a3 = LUT(wb, sel1);
      }
#pragma unroll
      for (int nb = 0; nb < NB; ++nb) mma16816(pr[nb], a0, a1, a2, a3, xb[nb][2 * j], xb[nb][2 * j + 1]);
    }
    if constexpr (INT4) {
      const __half2 ha = *reinterpret_cast<const __half2*>(&sc.x), hb = *reinterpret_cast<const __half2*>(&sc.y);
      const float sa = __low2float(ha), za = __high2float(ha) + 128.f, sbv = __low2float(hb), zb = __high2float(hb) + 128.f;
#pragma unroll
      for (int nb = 0; nb < NB; ++nb) {
        if constexpr (MODE == 0) {
          const float xs0 = XREG ? xsreg[XREG ? xi : 0] : xsum_s[kb];
          acc[0][0] = fmaf(sa, pr[0][0] - za * xs0, acc[0][0]);
          acc[0][2] = fmaf(sbv, pr[0][2] - zb * xs0, acc[0][2]);
        } else {
          const int c0 = 2 * t + 8 * nb;
          const float xs0 = c0 < B ? xsum_s[c0 * p.nkb + kb] : 0.f;
          const float xs1 = c0 + 1 < B ? xsum_s[(c0 + 1) * p.nkb + kb] : 0.f;
          acc[nb][0] = fmaf(sa, pr[nb][0] - za * xs0, acc[nb][0]);
          acc[nb][1] = fmaf(sa, pr[nb][1] - za * xs1, acc[nb][1]);
          acc[nb][2] = fmaf(sbv, pr[nb][2] - zb * xs0, acc[nb][2]);
          acc[nb][3] = fmaf(sbv, pr[nb][3] - zb * xs1, acc[nb][3]);
        }
      }
    } else {
      const float sa = __uint_as_float(sc.x), sbv = __uint_as_float(sc.y);
#pragma unroll
      for (int nb = 0; nb < NB; ++nb) {
        acc[nb][0] = fmaf(sa, pr[nb][0], acc[nb][0]);
        acc[nb][2] = fmaf(sbv, pr[nb][2], acc[nb][2]);
        if constexpr (MODE != 0) {
          acc[nb][1] = fmaf(sa, pr[nb][1], acc[nb][1]);
          acc[nb][3] = fmaf(sbv, pr[nb][3], acc[nb][3]);
        }
      }
    }
  };

  int jf = 0;            // tiles flushed so far (reduction-buffer sequence, same order as the reducer's)
  int ct = t_first;      // flat tile the accumulators belong to
  auto flush = [&]() {   // hand the partial sums of tile `ct` to the reducer (zeros if this warp had no unit of it)
    const int buf = jf % NBUF;
    if (jf >= NBUF) mbar_wait(rempty_a + buf * 8, (uint32_t)((jf / NBUF) - 1) & 1u);
    const int rbuf = buf * (kV4CW * REDW) + warp * REDW;
    if constexpr (MODE == 0) {
      if (t == 0) { red[rbuf + g] = acc[0][0]; red[rbuf + g + 8] = acc[0][2]; }
    } else {
#pragma unroll
      for (int nb = 0; nb < NB; ++nb) {
        *reinterpret_cast<float2*>(red_at(rbuf + g * COLS + 8 * nb + 2 * t)) = make_float2(acc[nb][0], acc[nb][1]);
        *reinterpret_cast<float2*>(red_at(rbuf + (g + 8) * COLS + 8 * nb + 2 * t)) = make_float2(acc[nb][2], acc[nb][3]);
      }
    }
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
#pragma unroll
      for (int i = 0; i < 4; ++i) acc[nb][i] = 0.f;
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(rfull_a + buf * 8);
    ++jf; ++ct;
  };

  if constexpr (WIN) {
    const int ntl = t_last - t_first + 1;
    int s = 0, q_now = 0;        // ring slot; flush sequence (window * ntl + tile) the accumulators belong to
    uint32_t ph = 0;
    auto flush_to = [&](int q) { while (q_now < q) { flush(); ++q_now; } };
    for (int w = 0; w < p.nwin; ++w) {
      if (w > 0) asm volatile("bar.sync 1, %0;" ::"n"(kV4Consumers) : "memory");   // every warp is done with the previous window's rows
      stage_window(w);
      const int wl = min(p.wu, upt - w * p.wu), tot = ntl * wl;
      for (int j0 = 0; j0 < tot; j0 += kV4CW) {
        const int j = j0 + warp;
        mbar_wait(full_a + s * 8, ph);
        if (j < tot) {
          const int tl = j / wl, kuw = j - tl * wl;
          const uint32_t sp = stage_addr(s) + (uint32_t)warp * UB;
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
          flush_to(w * ntl + tl);
          if (p.debug != 1) {
#pragma unroll
            for (int r = 0; r < RU; ++r) {
#pragma unroll
              for (int b = 0; b < BPR; ++b) compute(wv[r][b], sc[r], (kuw * RU + r) * BPR + b, 0);
            }
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty_a + s * 8);
        if (++s == S) { s = 0; ph ^= 1u; }
      }
      // hand in every partial sum of this window BEFORE the window barrier: a warp that kept its sums until its next unit
      // would leave the reducer (and through the reduction buffers the other warps) waiting across that barrier -> deadlock
      flush_to((w + 1) * ntl);
    }
    if (tid == 0 && p.debug != 3) v4_stamp(p, 4);
    return;
  }
  int u = u_first;
  int tf = tf_first, ku = ku_first;
  int s = 0;
  uint32_t ph = 0;
  for (; u < u1; u += kV4CW) {
    mbar_wait(full_a + s * 8, ph);
    if (tid == 0 && u == u0) { if (p.debug == 3) v4_stamp_cyc(p, 4, c_rel); else v4_stamp(p, 3); }
    const uint32_t sp = stage_addr(s) + (uint32_t)warp * UB;
    uint4 w[RU][BPR];
    uint2 sc[RU];
#pragma unroll
    for (int r = 0; r < RU; ++r) {
#pragma unroll
      for (int b = 0; b < BPR; ++b) {
        const uint32_t wa = sp + r * REC + b * 512 + lane * 16;
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(w[r][b].x), "=r"(w[r][b].y), "=r"(w[r][b].z), "=r"(w[r][b].w) : "r"(wa));
      }
      asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(sc[r].x), "=r"(sc[r].y) : "r"(sp + r * REC + BPR * 512 + g * 8));
    }
    while (ct < tf) flush();
    if (p.debug != 1) {
#pragma unroll
      for (int r = 0; r < RU; ++r) {
#pragma unroll
        for (int b = 0; b < BPR; ++b) compute(w[r][b], sc[r], (ku * RU + r) * BPR + b, r * BPR + b);
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(empty_a + s * 8);
    ku += kV4CW;
    while (ku >= upt) { ku -= upt; ++tf; }
    if (++s == S) { s = 0; ph ^= 1u; }
  }
  while (ct <= t_last) flush();
  if (tid == 0 && p.debug != 3) v4_stamp(p, 4);
  if constexpr (ATT) {
    // ---- fused decode attention: head `cta` on this CTA, att_nsplit key ranges on 128-thread sub-blocks ---------------
    asm volatile("bar.sync 1, %0;" ::"n"(kV4Consumers) : "memory");   // every consumer warp is done with the ring: reuse it
    const int sub = warp >> 2;
    if (cta < p.att.n_head && sub < p.att_nsplit) {
      const int t128 = tid & 127;
      if (t128 == 0) {   // the head's q, k and v rows (att_tph row tiles, stored by up to a few CTAs' reducers)
        const long long t0 = clock64();
        while (ld_acquire_gpu(p.att_cnt + cta) < (uint32_t)p.att_tph) {
          if (clock64() - t0 > 4000000000LL) { *reinterpret_cast<volatile uint32_t*>(p.ws) = 3u; break; }   // error word, no hang
        }
      }
      struct SubSync {
        int id;
        __device__ __forceinline__ void operator()() const { asm volatile("bar.sync %0, 128;" ::"r"(id) : "memory"); }
      };
      struct AllSync {
        int n;
        __device__ __forceinline__ void operator()() const { asm volatile("bar.sync 6, %0;" ::"r"(n) : "memory"); }
      };
      const SubSync ssync{2 + sub};
      ssync();
      AttnSmem<128, 1>* smA = reinterpret_cast<AttnSmem<128, 1>*>(sb + pl.ringA);
      if (t128 == 0 && sub == 0) v4_stamp(p, 6);   // trace: counter wait over
      attn_decode_item<128, 1, 16, SubSync, false, true, AllSync, 2>(p.att, smA[sub], t128, sub, cta, 0, ssync, AllSync{p.att_nsplit * 128});
      if (t128 == 0 && sub == 0) v4_stamp(p, 7);   // trace: attention done
    }
  }
}

static unsigned long long* g_v4_trace = nullptr;   // development: device buffer [slots][kV4MaxCtas][8]
static int g_v4_trace_slots = 0, g_v4_trace_next = 0;
unsigned long long* qgemv_v4_take_trace_slot() {
  return (g_v4_trace && g_v4_trace_next < g_v4_trace_slots) ? g_v4_trace + (size_t)(g_v4_trace_next++) * kV4MaxCtas * 8 : nullptr;
}
void qgemv_v4_set_trace(void* buf, int slots) { g_v4_trace = reinterpret_cast<unsigned long long*>(buf); g_v4_trace_slots = slots; g_v4_trace_next = 0; }

static std::atomic<int> g_v4_occ2{0};     // experiment: int4 batch-1 launches as two co-resident CTAs per SM
static std::atomic<int> g_v4_kvhint{1};   // K / V cache L2 hint honoured (tune knob 14 switches it off)
static std::atomic<int> g_v4_split{-1};   // -1 auto, 0 never split entities, 1 always
static std::atomic<int> g_v4_xreg{1};     // activation fragments in registers where the shape allows
void qgemv_v4_tune(int split, int xreg) { g_v4_split = split; g_v4_xreg = xreg; }
void qgemv_v4_tune_occ2(int on) { g_v4_occ2 = on; }
void qgemv_v4_tune_kvhint(int on) { g_v4_kvhint = on; }

template <int RU, bool INT4>
static cudaError_t launch_v4_occ2(const V4Params& p, int gx, cudaStream_t st) {
  static PerDeviceOnce once;
  if (once.need()) {
    cudaError_t e = cudaFuncSetAttribute(qgemv_v4_kernel<RU, INT4, 0, false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kV4DynSmem2);
    if (e != cudaSuccess) return e;
    once.done();
  }
  return launch(qgemv_v4_kernel<RU, INT4, 0, false, 2>, dim3(gx), dim3(kV4Threads), (size_t)kV4DynSmem2, st, p);
}

template <int RU, bool INT4, int MODE, bool XREG>
static cudaError_t launch_v4(const V4Params& p, int gx, cudaStream_t st) {
  static PerDeviceOnce once;  // per instantiation
  if (once.need()) {
    cudaError_t e = cudaFuncSetAttribute(qgemv_v4_kernel<RU, INT4, MODE, XREG>, cudaFuncAttributeMaxDynamicSharedMemorySize, kV4DynSmem);
    if (e != cudaSuccess) return e;
    once.done();
  }
  return launch(qgemv_v4_kernel<RU, INT4, MODE, XREG>, dim3(gx), dim3(kV4Threads), (size_t)kV4DynSmem, st, p);
}

template <int RU, bool INT4, int MODE>
static cudaError_t launch_v4_win(const V4Params& p, int gx, cudaStream_t st) {
  static PerDeviceOnce once;  // per instantiation
  if (once.need()) {
    cudaError_t e = cudaFuncSetAttribute(qgemv_v4_kernel<RU, INT4, MODE, false, 1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kV4DynSmem);
    if (e != cudaSuccess) return e;
    once.done();
  }
  return launch(qgemv_v4_kernel<RU, INT4, MODE, false, 1, true>, dim3(gx), dim3(kV4Threads), (size_t)kV4DynSmem, st, p);
}

template <int RU, bool INT4, bool XREG>
static cudaError_t launch_v4_att(const V4Params& p, int gx, cudaStream_t st) {
  static PerDeviceOnce once;  // per instantiation
  if (once.need()) {
    cudaError_t e = cudaFuncSetAttribute(qgemv_v4_kernel<RU, INT4, 0, XREG, 1, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kV4DynSmem);
    if (e != cudaSuccess) return e;
    once.done();
  }
  return launch(qgemv_v4_kernel<RU, INT4, 0, XREG, 1, false, true>, dim3(gx), dim3(kV4Threads), (size_t)kV4DynSmem, st, p);
}

template <bool INT4, int MODE>
static cudaError_t launch_v4_ru(int ru, const V4Params& p, int gx, cudaStream_t st) {
  constexpr int RUMAX = INT4 ? 2 : 4;
  if constexpr (MODE == 0) {
    if (p.att_nsplit > 0) {   // the host checked ru == RUMAX
      return (p.upt == kV4CW && g_v4_xreg.load()) ? launch_v4_att<RUMAX, INT4, true>(p, gx, st) : launch_v4_att<RUMAX, INT4, false>(p, gx, st);
    }
  }
  if constexpr (MODE != 0) {
    if (p.nwin > 1) return ru == RUMAX ? launch_v4_win<RUMAX, INT4, MODE>(p, gx, st) : launch_v4_win<1, INT4, MODE>(p, gx, st);
  }
  if (ru == RUMAX) {
    if constexpr (MODE == 0 && INT4) {
      if (g_v4_occ2.load()) return launch_v4_occ2<RUMAX, true>(p, gx, st);
    }
    if constexpr (MODE == 0) {
      if (p.upt == kV4CW && g_v4_xreg.load()) return launch_v4<RUMAX, INT4, 0, true>(p, gx, st);
    }
    return launch_v4<RUMAX, INT4, MODE, false>(p, gx, st);
  }
  if constexpr (!INT4) {
    if (ru == 2) return launch_v4<2, false, MODE, false>(p, gx, st);
  }
  return launch_v4<1, INT4, MODE, false>(p, gx, st);
}

size_t qgemv_v4_workspace_bytes() { return kV4WsData + (size_t)kV4MaxCtas * kV4WsPerCta; }

int qgemv_v4_launch(const b200_qgemv_tiled_args* a0, int debug, cudaStream_t st, bool dry) {
  // row folding: run the kernel on the (N / F, F * K) view with F activation rows per true row (zero outside their segment)
  const int fold = a0->fold > 1 ? a0->fold : 1;
  b200_qgemv_tiled_args av = *a0;
  av.N = a0->N / fold; av.K = a0->K * fold; av.batch = a0->batch * fold;
  const b200_qgemv_tiled_args* a = &av;
  const bool int4 = a->qtype == B200_Q_INT4;
  const bool swiglu = a->epilogue == B200_EPI_SWIGLU;
  const int nkb = (int)(a->K / 64);
  const int nrec = int4 ? nkb / 2 : nkb;                 // records per tile
  const int mode = a->batch == 1 ? 0 : (a->batch <= 8 ? 1 : 2);
  const int cols = mode == 2 ? 16 : 8;
  const bool gapred = !int4 && mode != 0;                // batched nf4 / fp4: the reduction buffers sit in the table-row gaps
  int nbuf = mode == 0 ? 4 : 2;
  int red_bytes = gapred ? 0 : (mode == 0 ? 4 * kV4CW * 16 : nbuf * kV4CW * 16 * cols) * 4;
  const int xstride = (int)a->K + 8;
  const int gap = (!int4 && mode == 0 && nkb <= 256) ? 1 : 0;
  if (!dry && (a->workspace == nullptr || a->workspace_bytes < qgemv_v4_workspace_bytes())) return B200_ERR_UNSUPPORTED;
  // largest unit (records per unit) that divides the tile and leaves at least two ring stages
  int ru = 0;
  for (int cand = int4 ? 2 : 4; cand >= 1; cand >>= 1) {
    if (nrec % cand) continue;
    const int stage = kV4CW * cand * (int4 ? 1088 : 576);
    const V4Plan pl = v4_plan(1024u, stage, red_bytes, int4, a->batch, nkb, xstride, gap, (int4 && mode == 0 && g_v4_occ2.load()) ? kV4DynSmem2 : kV4DynSmem);
    if (pl.S >= 2) { ru = cand; break; }
  }
  if (ru == 0 && mode != 0 && !gapred) {
    // one reduction buffer instead of two when that lets the full activation rows stay resident (no k-windows)
    const int red1 = kV4CW * 16 * cols * 4;
    for (int cand = int4 ? 2 : 4; cand >= 1; cand >>= 1) {
      if (nrec % cand) continue;
      const V4Plan pl = v4_plan(1024u, kV4CW * cand * (int4 ? 1088 : 576), red1, int4, a->batch, nkb, xstride, gap, kV4DynSmem);
      if (pl.S >= 2) { ru = cand; nbuf = 1; red_bytes = red1; break; }
    }
  }
  const int64_t ntiles = (a->N + 15) / 16;
  const int ts = swiglu ? 2 : 1;
  // k-windows: batch > 1 whose full activation rows leave no room for the ring.  The rows are staged one window of `wu`
  // units at a time; the CTA sweeps its tiles once per window and the reducer keeps per-tile partial sums.
  int wu = 0, nwin = 1, win_nkb = 0, win_xstride = 0;
  if (ru == 0 && mode != 0 && fold == 1 && !(a->tp && a->tp->slots)) {
    int sms = sm_count();
    if (sms > kV4MaxCtas) sms = kV4MaxCtas;
    const int64_t gxw = ntiles < sms ? ntiles : sms;
    const int extra = (int)(((ntiles + gxw - 1) / gxw) * ts) * 16 * cols * 4;
    // fewest windows first (each costs a barrier pair, a reload of the rows and a round of partial sums): the largest window
    // that leaves two ring stages; the windows are then equalised, which usually frees room for a third stage
    for (int cand = int4 ? 2 : 4; cand >= 1 && ru == 0; cand = cand > 1 ? 1 : 0) {   // the largest unit, then single records
      if (nrec % cand) continue;
      const int upt_c = nrec / cand, uk = cand * (int4 ? 128 : 64), stage = kV4CW * cand * (int4 ? 1088 : 576);
      for (int w = upt_c - 1; w >= 1; --w) {
        const V4Plan pl = v4_plan(1024u, stage, red_bytes, int4, a->batch, w * uk / 64, w * uk + 8, 0, kV4DynSmem, extra);
        if (pl.S >= 2) {
          nwin = (upt_c + w - 1) / w;
          wu = (upt_c + nwin - 1) / nwin;      // equal windows
          ru = cand; win_nkb = wu * uk / 64; win_xstride = wu * uk + 8;
          break;
        }
      }
    }
  }
  if (ru == 0) return B200_ERR_UNSUPPORTED;
  const int64_t U = ntiles * ts * (nrec / ru);
  if (U <= 0 || U > 0x3fffffff) return B200_ERR_UNSUPPORTED;
  // fused decode attention: batch 1, whole row tiles per head, one CTA per head
  AttnParams att{};
  int att_nsplit = 0, att_tph = 0;
  if (a0->attn) {
    const b200_attn_decode_args* at = a0->attn;
    int sms = sm_count();
    if (sms > kV4MaxCtas) sms = kV4MaxCtas;
    if (fold != 1 || mode != 0 || swiglu || a0->epilogue != B200_EPI_STORE || a0->tp || nwin > 1 || ru != (int4 ? 2 : 4) ||
        (!dry && (at->qkv != a0->y || !a0->y)) || at->n_head > sms || ntiles < sms ||
        a0->N != (int64_t)3 * at->n_head * at->head_size || attn_fused_params(at, 4, &att) != B200_OK)
      return B200_ERR_UNSUPPORTED;
    att_nsplit = 4; att_tph = 3 * at->head_size / 16;
  }
  if (dry) return B200_OK;

  V4Params p{};
  p.tA = reinterpret_cast<const uint8_t*>(a->tiled);
  p.tB = swiglu ? reinterpret_cast<const uint8_t*>(a->tiled2) : p.tA;
  p.N = a->N; p.K = a->K; p.nkb = nkb; p.upt = nrec / ru; p.ts = ts; p.U = (int)U;
  p.qtype = a->qtype; p.batch = a->batch; p.epilogue = a->epilogue; p.debug = debug;
  p.x = reinterpret_cast<const __nv_bfloat16*>(a->x);
  p.delta = reinterpret_cast<const __nv_bfloat16*>(a->delta);
  p.x_out = reinterpret_cast<__nv_bfloat16*>(a->x_out);
  p.norm_w = reinterpret_cast<const __nv_bfloat16*>(a->norm_weight);
  p.eps = a->norm_eps;
  p.bias = reinterpret_cast<const __nv_bfloat16*>(a->bias);
  p.bias2 = reinterpret_cast<const __nv_bfloat16*>(a->bias2);
  p.residual = reinterpret_cast<const __nv_bfloat16*>(a->residual);
  p.y = reinterpret_cast<__nv_bfloat16*>(a->y);
  p.xstride = xstride; p.gap = gap;
  p.wu = wu; p.nwin = nwin; p.nbuf = nbuf;
  p.att = att; p.att_nsplit = att_nsplit; p.att_tph = att_tph;
  if (att_nsplit) { p.att_cnt = reinterpret_cast<unsigned int*>(reinterpret_cast<uint8_t*>(a->workspace) + 1024); p.att.fuse_cnt = p.att_cnt; }
  if (nwin > 1) { p.nkb = win_nkb; p.xstride = win_xstride; }
  p.fold = fold; p.Nout = a0->N; p.Kin = (int)a0->K;
  if (a0->kv_hint && a0->kv_hint->k && a0->kv_hint->v && a0->kv_hint->pos && g_v4_kvhint.load()) {
    const b200_kv_hint* h = a0->kv_hint;
    p.kvh_k = reinterpret_cast<const uint8_t*>(h->k); p.kvh_v = reinterpret_cast<const uint8_t*>(h->v); p.kvh_pos = h->pos;
    p.kvh_i64 = h->pos_is_int64; p.kvh_chunks = h->chunks; p.kvh_row = h->row_bytes; p.kvh_max = h->max_rows; p.kvh_stride = h->chunk_stride;
  }
  p.ws = reinterpret_cast<uint8_t*>(a->workspace);
  if (a->tp) {
    const b200_tp_handoff* h = a->tp;
    p.tp_world = h->world;
    if (h->send_seq) {
      for (int r = 0; r < h->world; ++r) p.tp_peer_y[r] = reinterpret_cast<uint32_t*>(h->peer_y[r]);
      p.tp_send_seq = reinterpret_cast<const uint32_t*>(h->send_seq);
    }
    if (h->slots) {
      p.tp_slots = reinterpret_cast<const uint32_t*>(h->slots);
      p.tp_slot_stride = h->slot_stride / 4;
      p.tp_recv_seq = reinterpret_cast<uint32_t*>(h->recv_seq);
      p.tp_recv_done = reinterpret_cast<uint32_t*>(h->recv_done);
      p.tp_reduced = reinterpret_cast<uint32_t*>(h->reduced);
    }
  }
  p.trace = qgemv_v4_take_trace_slot();
  int gx = sm_count();
  if (gx > kV4MaxCtas) gx = kV4MaxCtas;
  // Whole entities per CTA unless the resulting imbalance costs more than the partial hand-off of a split entity
  // (~2 us exposed: fence + flag + two L2 round trips): bytes of imbalance at ~6 TB/s.
  const int64_t E = ntiles;                       // entities (tiles, or gate/up pairs)
  const int64_t ent_bytes = (int64_t)ts * nrec * (int4 ? 1088 : 576);
  const int force = (nwin > 1 || att_nsplit) ? 0 : g_v4_split.load();   // k-windows and the fused attention keep whole entities per CTA
  if (E >= gx) {
    const double waste_us = (double)(((E + gx - 1) / gx) * gx - E) / gx * ent_bytes / 6.0e6;   // idle share of the last wave
    p.split = force >= 0 ? force : (waste_us > 2.5 ? 1 : 0);
  } else {
    p.split = force >= 0 ? force : ((double)ent_bytes / 6.0e6 > 4.0 ? 1 : 0);                   // fewer entities than SMs
    if (!p.split) gx = (int)E;
  }
  if ((int64_t)gx > U) gx = (int)U;
  cudaError_t e;
  if (int4) e = mode == 0 ? launch_v4_ru<true, 0>(ru, p, gx, st) : mode == 1 ? launch_v4_ru<true, 1>(ru, p, gx, st) : launch_v4_ru<true, 2>(ru, p, gx, st);
  else e = mode == 0 ? launch_v4_ru<false, 0>(ru, p, gx, st) : mode == 1 ? launch_v4_ru<false, 1>(ru, p, gx, st) : launch_v4_ru<false, 2>(ru, p, gx, st);
  if (e != cudaSuccess) {
    set_error("b200_qgemv_tiled: launch failed: %s", cudaGetErrorString(e));
    return B200_ERR_CUDA;
  }
  return B200_OK;
}

}  // namespace b200
