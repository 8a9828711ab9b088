// Persistent decode-step kernel: a whole T = 1 forward as ONE launch.
//
// Replaces the per-token chain generate/base.py:44-47 -> lit_gpt/model.py:499-519 -> Block.forward 572-593 (x n_layer).
// The step is a PROGRAM of ops (b200_step_op: GEMV / ATTN / EMBED) over caller-owned device buffers.  One persistent
// CTA per SM (16 consumer warps + 1 producer warp) interprets it:
//   * the PRODUCER warp walks the program on its own and streams the tiled weights of every GEMV op, in program
//     order, through a ring of shared-memory stages (TMA bulk copies on full/empty mbarriers).  It is throttled only
//     by ring space, never by op boundaries, so while the consumers sit in a grid barrier, an RMSNorm prologue or the
//     attention op, the weights of the NEXT matrix are already arriving: HBM does not idle between ops, which is
//     what a chain of per-op launches cannot do;
//   * the CONSUMER warps run each op (GEMV: same dequant + mma.sync inner loop as qgemv_stream.cu; ATTN: the
//     attn_body.cuh item body on four 128-thread sub-blocks; EMBED: row gather) and meet in a grid-wide barrier
//     (one atomic counter, release/acquire at gpu scope) after every op;
//   * the 64 KB dequant table, the mbarriers and the ring are set up once per step instead of once per matrix.
// Cross-CTA activations are read with ld.global.cg (L2) so no stale L1 line can be observed inside the kernel.
#include "attn_body.cuh"
#include "stream_common.cuh"

namespace b200 {

struct StepParams {
  const b200_step_op* ops;
  int n_ops;
  int batch;
  int qtype;
  int max_k;            // largest K of any GEMV op (sizes the activation rows in shared memory)
  int xstride;          // max_k + 8
  int red_bufs;         // 1 or 2 reduction buffers
  int red_cols;         // batch columns kept in the reduction buffers (2 when batch <= 2, else 8 * NB)
  int scratch_bytes;    // reduction buffers / attention scratch region
  int attn_nsub;        // attention sub-blocks (128 threads each) per CTA, 1..4
  unsigned int* sync;   // [0] barrier counter (zeroed before the launch), [1] error flag
  int lookahead_kb;     // L2 look-ahead of the weight stream per CTA (0 = off)
  long long* trace;     // optional: 8 clock64() slots per op, written by CTA 0
  int debug;            // experiments: bit 0 = consumers skip the dequant/MMA work, bit 1 = producer splits each tile copy in 4
  unsigned int zero;    // always 0; lets an address depend on loaded registers without the compiler folding it away
};

constexpr int kDW = 16;                          // consumer warps (they split K)
constexpr int kDThreadsAll = (kDW + 1) * 32;     // + 1 producer warp
constexpr int kDConsumers = kDW * 32;
constexpr int kDMaxStages = 8;
constexpr int kDDynSmem = 227 * 1024;

__device__ __forceinline__ uint4 ldcg_v4(const void* p) { return __ldcg(reinterpret_cast<const uint4*>(p)); }

// Walks the GEMV ops of the program and yields, for this CTA, the (tile0, tile1, round) stages in consumption
// order.  Producer and consumers (and the L2 look-ahead cursor) run identical copies of it.
template <bool INT4>
struct StageWalk {
  static constexpr int KPU = INT4 ? 2 : 1;
  const b200_step_op* ops;
  int n_ops;
  int oi;          // current op index (may be a non-GEMV op: skipped by next_op)
  // current GEMV op
  const uint8_t* tA;
  const uint8_t* tB;
  int nku, nrounds, u, u1, ustep, swiglu, r;
  __device__ __forceinline__ void init(const b200_step_op* o, int n) { ops = o; n_ops = n; oi = -1; u = u1 = 0; r = 0; nrounds = 1; }
  // position on op `i` (must be a GEMV op)
  __device__ __forceinline__ void enter(int i) {
    const b200_step_op& op = ops[i];
    oi = i;
    swiglu = op.epilogue == B200_EPI_SWIGLU;
    tA = reinterpret_cast<const uint8_t*>(op.w);
    tB = swiglu ? reinterpret_cast<const uint8_t*>(op.w2) : tA;
    const int nkb = (int)(op.K >> 6);
    nku = nkb / KPU;
    nrounds = (nku + kDW - 1) / kDW;
    const int nunits = (int)((op.N + 15) >> 4);
    u = (int)(((int64_t)nunits * blockIdx.x) / gridDim.x);
    u1 = (int)(((int64_t)nunits * (blockIdx.x + 1)) / gridDim.x);
    ustep = swiglu ? 1 : 2;
    r = 0;
  }
  __device__ __forceinline__ bool op_done() const { return u >= u1; }
  __device__ __forceinline__ void advance() {
    if (++r == nrounds) { r = 0; u += ustep; }
  }
  // advance to the next GEMV op that has work for this CTA; false at the end of the program
  __device__ __forceinline__ bool next_gemv() {
    for (int i = oi + 1; i < n_ops; ++i) {
      if (ops[i].kind != B200_OP_GEMV) continue;
      enter(i);
      if (!op_done()) return true;
    }
    oi = n_ops;
    return false;
  }
};

template <int NB, bool INT4>
__global__ void __launch_bounds__(kDThreadsAll, 1) decode_step_kernel(const StepParams sp) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  constexpr int COLS = 8 * NB;
  constexpr int REC = INT4 ? 1088 : 576;         // one k-unit (NF4: 64 k, INT4: 128 k) of one tile
  constexpr int KPU = INT4 ? 2 : 1;              // k-blocks per unit
  constexpr int STAGE_TILE = kDW * REC;          // 16 consecutive units of one tile
  constexpr int STAGE = 2 * STAGE_TILE;          // two tiles
  const uint32_t b0 = smem_u32(smem_raw);
  uint8_t* const sb = smem_raw - b0;             // generic pointer to shared-window address 0

  // ---- shared-memory plan by ABSOLUTE shared-window addresses (the table must sit at a 64 KB aligned one) --------
  uint32_t off = (b0 + 127u) & ~127u;
  const uint32_t bars_a = off;          off += 2 * kDMaxStages * 8;                     // full[6], empty[6]
  off = (off + 127u) & ~127u;
  const uint32_t op_a = off;            off += 256;                                     // the current op (copied from global once per op)
  const uint32_t red_a = off;           off += (uint32_t)sp.scratch_bytes;              // reduction buffers / attention scratch
  const uint32_t ssq_a = off;           off += kDW * 16 * 4;                            // [16 warps][16]
  const uint32_t rstd_a = off;          off += 16 * 4;
  const uint32_t xsum_a = off;          off += INT4 ? (((uint32_t)(sp.batch * (sp.max_k >> 6)) + 3u) & ~3u) * 4 : 0;
  const uint32_t xs_a = off;            off += (((uint32_t)sp.batch * sp.xstride * 2) + 127u) & ~127u;
  const uint32_t ringA_a = off;
  // int4 dequantises arithmetically: no table, the whole remainder is ring
  const uint32_t lut_a = INT4 ? off : (off + 0xFFFFu) & ~0xFFFFu;                       // 64 KB aligned
  const uint32_t ringC_a = INT4 ? off : lut_a + 0x10000u;
  const uint32_t end_a = b0 + kDDynSmem;
  const int nA = (int)((lut_a - ringA_a) / STAGE);
  const int nC = ringC_a < end_a ? (int)((end_a - ringC_a) / STAGE) : 0;
  const int S = min(kDMaxStages, nA + nC);
  if (S < 2 || ringC_a > end_a) __trap();        // the host sized the launch so that this cannot happen
  auto stage_addr = [&](int s) -> uint32_t { return s < nC ? ringC_a + s * STAGE : ringA_a + (s - nC) * STAGE; };

  float* red = reinterpret_cast<float*>(sb + red_a);
  float* s_ssq = reinterpret_cast<float*>(sb + ssq_a);
  float* s_rstd = reinterpret_cast<float*>(sb + rstd_a);
  float* xsum_s = reinterpret_cast<float*>(sb + xsum_a);
  __nv_bfloat16* x_s = reinterpret_cast<__nv_bfloat16*>(sb + xs_a);
  uint32_t* lut = reinterpret_cast<uint32_t*>(sb + lut_a);
  static_assert(sizeof(b200_step_op) <= 256 && sizeof(b200_step_op) % 4 == 0, "op slot");
  constexpr int OP_WORDS = (int)(sizeof(b200_step_op) / 4);
  uint32_t* s_opw = reinterpret_cast<uint32_t*>(sb + op_a);
  // the op being executed, in shared memory: fields are re-read at their use sites instead of living in registers
  const b200_step_op& op = *reinterpret_cast<const b200_step_op*>(sb + op_a);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int B = sp.batch;
  const int RC = sp.red_cols;                    // columns per row in the reduction buffers
  const int RED_ONE = kDW * 2 * 16 * RC;         // floats per reduction buffer

  if (tid == 0) {
    for (int s = 0; s < kDMaxStages; ++s) { mbar_init(bars_a + s * 8, 1); mbar_init(bars_a + (kDMaxStages + s) * 8, kDW); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  if (warp == kDW) {
    // ================================ PRODUCER WARP (one elected lane) ==========================================
    if (lane == 0) {
      const uint64_t pol = make_evict_first_policy();
      StageWalk<INT4> w, pf;
      w.init(sp.ops, sp.n_ops);
      pf.init(sp.ops, sp.n_ops);
      int64_t ahead = 0;                              // bytes the look-ahead cursor is in front of the load cursor
      const int64_t want = (int64_t)sp.lookahead_kb << 10;
      bool pf_live = want > 0 && pf.next_gemv();
      int n = 0, ps = 0;          // global stage counter, ring slot
      uint32_t pph = 1;           // parity to wait for on the slot's EMPTY barrier (first lap: no wait)
      while (w.next_gemv()) {
        while (!w.op_done()) {
          const bool has1 = w.swiglu || w.u + 1 < w.u1;
          const int nu = min(kDW, w.nku - w.r * kDW);
          const uint32_t bytes = (uint32_t)nu * REC;
          // L2 look-ahead: keep `want` bytes of this CTA's future stages requested from HBM
          while (pf_live && ahead < want) {
            const bool ph1 = pf.swiglu || pf.u + 1 < pf.u1;
            const uint32_t pb = (uint32_t)min(kDW, pf.nku - pf.r * kDW) * REC;
            prefetch_l2_bulk(pf.tA + ((int64_t)pf.u * pf.nku + pf.r * kDW) * REC, pb);
            if (ph1) prefetch_l2_bulk(pf.tB + ((int64_t)(pf.swiglu ? pf.u : pf.u + 1) * pf.nku + pf.r * kDW) * REC, pb);
            ahead += ph1 ? 2 * pb : pb;
            pf.advance();
            if (pf.op_done()) pf_live = pf.next_gemv();
          }
          if (n >= S) mbar_wait(bars_a + (kDMaxStages + ps) * 8, pph);
          const uint32_t bar = bars_a + ps * 8, dst = stage_addr(ps);
          mbar_expect_tx(bar, has1 ? 2 * bytes : bytes);
          const uint8_t* srcA = w.tA + ((int64_t)w.u * w.nku + w.r * kDW) * REC;
          const uint8_t* srcB = w.tB + ((int64_t)(w.swiglu ? w.u : w.u + 1) * w.nku + w.r * kDW) * REC;
          if ((sp.debug & 2) && nu == kDW) {
            constexpr uint32_t Q = STAGE_TILE / 4;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              bulk_g2s(dst + q * Q, srcA + q * Q, Q, bar, pol);
              if (has1) bulk_g2s(dst + STAGE_TILE + q * Q, srcB + q * Q, Q, bar, pol);
            }
          } else {
            bulk_g2s(dst, srcA, bytes, bar, pol);
            if (has1) bulk_g2s(dst + STAGE_TILE, srcB, bytes, bar, pol);
          }
          ahead -= has1 ? 2 * bytes : bytes;
          ++n;
          if (++ps == S) { ps = 0; pph ^= 1u; }
          w.advance();
        }
      }
    }
    return;
  }
  // ================================== CONSUMER WARPS ===============================================================
  for (int i = tid; i < (INT4 ? 0 : 256 * 32); i += kDConsumers) {
    const int e = i >> 5, l = i & 31;
    float lo, hi;
    if (INT4) { lo = (float)(e & 15); hi = (float)(e >> 4); }                      // (unused: int4 has no table)
    else { lo = code_value(sp.qtype, e >> 4); hi = code_value(sp.qtype, e & 15); } // even element = HIGH nibble
    lut[e * 64 + l] = pack_bf16(lo, hi);
  }
  if (tid < OP_WORDS) s_opw[tid] = __ldg(reinterpret_cast<const uint32_t*>(sp.ops) + tid);
  asm volatile("bar.sync 1, %0;" ::"n"(kDConsumers) : "memory");

  // table address = L | byte << 8 | lane << 2, assembled by ONE PRMT (L is 64 KB aligned)
  const uint32_t lane_full = lut_a | ((uint32_t)lane << 2);
  auto LUT = [&](uint32_t word, uint32_t sel) -> uint32_t {
    uint32_t v;
    asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(__byte_perm(word, lane_full, sel)));
    return v;
  };

  // ---- grid-wide barrier between ops: one counter, release (fence + atomic) / acquire (polling load + fence) ------
  unsigned int sync_target = 0;
  auto grid_sync = [&](int next_op, int cur_op) {
    asm volatile("bar.sync 1, %0;" ::"n"(kDConsumers) : "memory");
    if (tid >= 32 && tid < 32 + OP_WORDS) s_opw[tid - 32] = __ldg(reinterpret_cast<const uint32_t*>(sp.ops + next_op) + (tid - 32));
    if (tid == 0) {
      sync_target += gridDim.x;
      __threadfence();
      atomicAdd(sp.sync, 1u);
      if (sp.trace && blockIdx.x == 0) sp.trace[cur_op * 8 + 3] = clock64();
      unsigned int v;
      int spins = 0;
      do {
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(sp.sync) : "memory");
        if (++spins > (1 << 22)) { sp.sync[1] = 0xDEADu; __threadfence(); __trap(); }   // a lost CTA must not hang the GPU
      } while (v < sync_target);
      __threadfence();
      if (sp.trace && blockIdx.x == 0) sp.trace[cur_op * 8 + 4] = clock64();
    }
    asm volatile("bar.sync 1, %0;" ::"n"(kDConsumers) : "memory");
  };

  // ring state of the consumers (persists across ops; mirrors the producer's stage order)
  int ls = 0;              // ring slot of the next stage to read
  uint32_t lph = 0;        // its FULL-barrier parity
  int rs = 0;              // ring slot to release next
  int it = 0;              // reduction-buffer parity

  auto stamp = [&](int oi, int slot) {
    if (sp.trace && blockIdx.x == 0 && tid == 0) sp.trace[oi * 8 + slot] = clock64();
  };
  for (int oi = 0; oi < sp.n_ops; ++oi) {
    const int kind = op.kind;
    stamp(oi, 0);
    if (kind == B200_OP_GEMV) {
      // ============================================ GEMV op =====================================================
      const int nkb = (int)(op.K >> 6);
      const int nku = nkb / KPU;
      const int nrounds = (nku + kDW - 1) / kDW;
      const int nunits = (int)((op.N + 15) >> 4);
      const bool swiglu = op.epilogue == B200_EPI_SWIGLU;
      const int u0 = (int)(((int64_t)nunits * blockIdx.x) / gridDim.x);
      const int u1 = (int)(((int64_t)nunits * (blockIdx.x + 1)) / gridDim.x);
      const int ustep = swiglu ? 1 : 2;
      const int xstride = sp.xstride;

      if (u0 < u1 || (op.x_out && blockIdx.x == 0)) {
        const int64_t K = op.K;
        const __nv_bfloat16* xg = reinterpret_cast<const __nv_bfloat16*>(op.x);
        const __nv_bfloat16* dg = reinterpret_cast<const __nv_bfloat16*>(op.delta);
        __nv_bfloat16* xo = reinterpret_cast<__nv_bfloat16*>(op.x_out);
        const __nv_bfloat16* nw = reinterpret_cast<const __nv_bfloat16*>(op.norm_weight);
        const float eps = op.norm_eps;
        // ---- activation prologue: (x + delta) -> optional RMSNorm -> bf16 rows in shared memory -----------------
        const bool designated = blockIdx.x == 0;
        const int chunks = (int)(K >> 3);
        for (int b = 0; b < B; ++b) {
          float ss = 0.f;
          for (int c = tid; c < chunks; c += kDConsumers) {
            uint4 v = ldcg_v4(xg + (int64_t)b * K + c * 8);
            if (dg) {
              const uint4 dv = ldcg_v4(dg + (int64_t)b * K + c * 8);
              v.x = pack_bf16(bf16lo(v.x) + bf16lo(dv.x), bf16hi(v.x) + bf16hi(dv.x));
              v.y = pack_bf16(bf16lo(v.y) + bf16lo(dv.y), bf16hi(v.y) + bf16hi(dv.y));
              v.z = pack_bf16(bf16lo(v.z) + bf16lo(dv.z), bf16hi(v.z) + bf16hi(dv.z));
              v.w = pack_bf16(bf16lo(v.w) + bf16lo(dv.w), bf16hi(v.w) + bf16hi(dv.w));
              if (xo && designated) *reinterpret_cast<uint4*>(xo + (int64_t)b * K + c * 8) = v;
            }
            *reinterpret_cast<uint4*>(x_s + b * xstride + c * 8) = v;
            const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) { const float a = bf16lo(w4[i]), c2 = bf16hi(w4[i]); ss = fmaf(a, a, ss); ss = fmaf(c2, c2, ss); }
          }
          if (nw) {
            ss = warp_sum(ss);
            if (lane == 0) s_ssq[warp * 16 + b] = ss;
          }
        }
        if (nw) {
          asm volatile("bar.sync 1, %0;" ::"n"(kDConsumers) : "memory");
          if (tid < B) {
            float tot = 0.f;
#pragma unroll
            for (int w = 0; w < kDW; ++w) tot += s_ssq[w * 16 + tid];
            s_rstd[tid] = rsqrtf(tot / (float)K + eps);
          }
          asm volatile("bar.sync 1, %0;" ::"n"(kDConsumers) : "memory");
          for (int b = 0; b < B; ++b) {
            const float rstd = s_rstd[b];
            for (int c = tid; c < chunks; c += kDConsumers) {  // same thread owns the same chunk as above
              uint4 v = *reinterpret_cast<uint4*>(x_s + b * xstride + c * 8);
              const uint4 n4 = *reinterpret_cast<const uint4*>(nw + c * 8);
              // weight * (x * rstd), fp32, one cast (rmsnorm.py:24-25)
              v.x = pack_bf16(bf16lo(n4.x) * (bf16lo(v.x) * rstd), bf16hi(n4.x) * (bf16hi(v.x) * rstd));
              v.y = pack_bf16(bf16lo(n4.y) * (bf16lo(v.y) * rstd), bf16hi(n4.y) * (bf16hi(v.y) * rstd));
              v.z = pack_bf16(bf16lo(n4.z) * (bf16lo(v.z) * rstd), bf16hi(n4.z) * (bf16hi(v.z) * rstd));
              v.w = pack_bf16(bf16lo(n4.w) * (bf16lo(v.w) * rstd), bf16hi(n4.w) * (bf16hi(v.w) * rstd));
              *reinterpret_cast<uint4*>(x_s + b * xstride + c * 8) = v;
            }
          }
        }
        asm volatile("bar.sync 1, %0;" ::"n"(kDConsumers) : "memory");  // x_s is published
        if (INT4) {       // per (batch row, k-block) activation sums for the zero-point term
          xsum_blocks<kDConsumers>(x_s, xstride, B, nkb, xsum_s, tid);
          asm volatile("bar.sync 1, %0;" ::"n"(kDConsumers) : "memory");
        }
      }

      stamp(oi, 1);
      float acc0[NB][4], acc1[NB][4];
#pragma unroll
      for (int nb = 0; nb < NB; ++nb) {
#pragma unroll
        for (int i = 0; i < 4; ++i) acc0[nb][i] = acc1[nb][i] = 0.f;
      }

      // one k-block of both tiles: w0 / w1 = this lane's 16 B {row g bytes, row g+8 bytes}; s0 / s1 = scale words.
      // Tile 1 is computed even when absent (stale stage bytes, result never stored): no branch in the hot block.
      auto compute = [&](const uint4 w0, const uint4 w1, const uint2 s0, const uint2 s1, int kb) {
        uint32_t xb[NB][8];
#pragma unroll
        for (int nb = 0; nb < NB; ++nb) {
          const int b = g + 8 * nb;
          if (b < B) {
            const uint4* px = reinterpret_cast<const uint4*>(x_s + b * xstride + kb * 64 + t * 16);
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
            const uint32_t m = 0x000F000Fu, c = 0x43004300u;
#pragma unroll
            for (int nb = 0; nb < NB; ++nb) {
              mma16816(p0[nb], ((ua >> sh) & m) | c, ((ub >> sh) & m) | c, ((ua >> (sh + 4)) & m) | c, ((ub >> (sh + 4)) & m) | c, xb[nb][2 * j], xb[nb][2 * j + 1]);
              mma16816(p1[nb], ((va >> sh) & m) | c, ((vb >> sh) & m) | c, ((va >> (sh + 4)) & m) | c, ((vb >> (sh + 4)) & m) | c, xb[nb][2 * j], xb[nb][2 * j + 1]);
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
          const float s0a = __low2float(h0a), z0a = __high2float(h0a) + 128.f, s0b = __low2float(h0b), z0b = __high2float(h0b) + 128.f;
          const float s1a = __low2float(h1a), z1a = __high2float(h1a) + 128.f, s1b = __low2float(h1b), z1b = __high2float(h1b) + 128.f;
#pragma unroll
          for (int nb = 0; nb < NB; ++nb) {
            const int c0 = 2 * t + 8 * nb;
            const float xs0 = c0 < B ? xsum_s[c0 * nkb + kb] : 0.f;
            const float xs1 = c0 + 1 < B ? xsum_s[(c0 + 1) * nkb + kb] : 0.f;
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
        const __nv_bfloat16* bias = reinterpret_cast<const __nv_bfloat16*>(op.bias);
        const int64_t N = op.N;
        float h = v + (bias ? __bfloat162float(bias[n]) : 0.f);
        if (op.epilogue == B200_EPI_RESIDUAL) h = round_bf16(h) + ld_act_bf16(reinterpret_cast<const __nv_bfloat16*>(op.residual) + (int64_t)c * N + n);
        reinterpret_cast<__nv_bfloat16*>(op.y)[(int64_t)c * N + n] = __float2bfloat16_rn(h);
      };

      // Register-level software pipeline over the flattened (step, round) sequence: the stage of round i+1 is read
      // into registers before round i is computed; a stage is handed back to the producer one round after its read.
      const int nsteps = (u1 - u0 + ustep - 1) / ustep;
      const int total = nsteps * nrounds;
      uint4 c0[KPU], c1[KPU], n0[KPU], n1[KPU];
      uint2 cs0, cs1, ns0, ns1;
      auto read_stage = [&](int r, uint4 (&a0)[KPU], uint4 (&a1)[KPU], uint2& b0v, uint2& b1v) {
        mbar_wait(bars_a + ls * 8, lph);
        const int ku = r * kDW + warp;
        if (ku < nku) {
          const uint32_t spa = stage_addr(ls) + warp * REC;
          const uint32_t sa = spa + KPU * 512 + g * 8;
#pragma unroll
          for (int b = 0; b < KPU; ++b) {
            const uint32_t wa = spa + b * 512 + lane * 16;
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

      int u = u0, r = 0;
      bool has1 = swiglu || u + 1 < u1;
      const int nku_c = (sp.debug & 1) ? 0 : nku;
      const bool tr = sp.trace != nullptr && blockIdx.x == 0 && tid == 0;
      long long t_wait = 0, t_comp = 0, t_red = 0, tc = tr ? clock64() : 0;
      auto lap = [&](long long& accu) {
        if (tr) { const long long now = clock64(); accu += now - tc; tc = now; }
      };
      if (total > 0) read_stage(0, c0, c1, cs0, cs1);
      lap(t_wait);
      for (int i = 0; i < total; ++i) {
        int nu_ = u, nr = r + 1;  // position of round i+1
        if (nr == nrounds) { nr = 0; nu_ = u + ustep; }
        const bool nhas1 = swiglu || nu_ + 1 < u1;
        if (i + 1 < total) read_stage(nr, n0, n1, ns0, ns1);
        lap(t_wait);
        const int ku = r * kDW + warp;
        if (ku < nku_c) {
#pragma unroll
          for (int b = 0; b < KPU; ++b) compute(c0[b], c1[b], cs0, cs1, ku * KPU + b);
        }
        __syncwarp();
        // Hand stages back to the producer as early as their bytes are in registers: round 0's stage after its
        // compute, and the stage read at the top of THIS iteration (round i+1) now -- its loads were issued a whole
        // compute ago; the barrier address is made to depend on one register of every load (sp.zero == 0), so the
        // arrive cannot be issued before they have landed.
        if (i == 0) {
          if (lane == 0) mbar_arrive(bars_a + (kDMaxStages + rs) * 8);
          if (++rs == S) rs = 0;
        }
        if (i + 1 < total) {
          uint32_t dep = ns0.x ^ ns1.x;
#pragma unroll
          for (int b = 0; b < KPU; ++b) dep ^= n0[b].x ^ n1[b].x;
          if (lane == 0) mbar_arrive(bars_a + (kDMaxStages + rs) * 8 + (dep & sp.zero));
          if (++rs == S) rs = 0;
        }
#pragma unroll
        for (int b = 0; b < KPU; ++b) { c0[b] = n0[b]; c1[b] = n1[b]; }
        cs0 = ns0; cs1 = ns1;
        lap(t_comp);
        if (nr == 0) {
          // ---- end of a step: reduce the 16 k-slices through shared memory ---------------------------------------
          float* rb = red + (sp.red_bufs == 2 ? (it & 1) : 0) * RED_ONE;
          float* rbuf = rb + warp * (2 * 16 * RC);
#pragma unroll
          for (int nb = 0; nb < NB; ++nb) {
            const int col = 8 * nb + 2 * t;
            if (col < RC) {
              *reinterpret_cast<float2*>(rbuf + g * RC + col) = make_float2(acc0[nb][0], acc0[nb][1]);
              *reinterpret_cast<float2*>(rbuf + (g + 8) * RC + col) = make_float2(acc0[nb][2], acc0[nb][3]);
              *reinterpret_cast<float2*>(rbuf + 16 * RC + g * RC + col) = make_float2(acc1[nb][0], acc1[nb][1]);
              *reinterpret_cast<float2*>(rbuf + 16 * RC + (g + 8) * RC + col) = make_float2(acc1[nb][2], acc1[nb][3]);
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) acc0[nb][k] = acc1[nb][k] = 0.f;
          }
          asm volatile("bar.sync 1, %0;" ::"n"(kDConsumers) : "memory");
          if (tid < 2 * 16 * RC) {  // one thread per (tile, row, batch column)
            const int tl = tid / (16 * RC), rr = (tid / RC) % 16, c = tid % RC;
            float v = 0.f;
#pragma unroll
            for (int w = 0; w < kDW; ++w) v += rb[w * (2 * 16 * RC) + tl * 16 * RC + rr * RC + c];
            if (swiglu) {
              // gate (tile 0) and up (tile 1) values of the same row meet in the gate thread through shared memory
              if (tl == 1) rb[16 * RC + rr * RC + c] = v;  // warp 0's tile-1 slot: only this thread ever reads it
              asm volatile("bar.sync 2, %0;" ::"r"(2 * 16 * RC) : "memory");
              if (tl == 0 && c < B) {
                const int64_t nrow = (int64_t)u * 16 + rr, N = op.N;
                const __nv_bfloat16* bias = reinterpret_cast<const __nv_bfloat16*>(op.bias);
                const __nv_bfloat16* bias2 = reinterpret_cast<const __nv_bfloat16*>(op.bias2);
                if (nrow < N) {
                  const float a = round_bf16(v + (bias ? __bfloat162float(bias[nrow]) : 0.f));
                  const float b = round_bf16(rb[16 * RC + rr * RC + c] + (bias2 ? __bfloat162float(bias2[nrow]) : 0.f));
                  reinterpret_cast<__nv_bfloat16*>(op.y)[(int64_t)c * N + nrow] = __float2bfloat16_rn(silu_bf16_s(a) * b);
                }
              }
            } else if (c < B && (tl == 0 || has1)) {
              const int64_t nrow = (int64_t)(u + tl) * 16 + rr;
              if (nrow < op.N) store_out(nrow, c, v);
            }
          }
          if (sp.red_bufs == 1) asm volatile("bar.sync 1, %0;" ::"n"(kDConsumers) : "memory");
          ++it;
          lap(t_red);
        }
        u = nu_; r = nr; has1 = nhas1;
      }
      if (tr) { sp.trace[oi * 8 + 5] = t_wait; sp.trace[oi * 8 + 6] = t_comp; sp.trace[oi * 8 + 7] = t_red; }
    } else if (kind == B200_OP_ATTN) {
      // ============================================ ATTN op =====================================================
      AttnParams ap;
      ap.qkv = reinterpret_cast<const __nv_bfloat16*>(op.x);
      ap.cos = op.cos; ap.sin = op.sin;
      ap.table_bf16 = op.table_dtype == B200_BF16; ap.table_presel = op.table_preselected; ap.rope_n = op.rope_n_elem;
      ap.pos = op.pos; ap.pos_i64 = op.pos_is_int64;
      ap.kc = reinterpret_cast<__nv_bfloat16*>(op.k_cache); ap.vc = reinterpret_cast<__nv_bfloat16*>(op.v_cache);
      ap.out = reinterpret_cast<__nv_bfloat16*>(op.y);
      ap.batch = B; ap.n_head = op.n_head; ap.n_groups = op.n_groups; ap.hs = op.head_size; ap.max_seq = op.max_seq;
      ap.qpk = op.n_head / op.n_groups;
      ap.nsplit = op.num_splits;
      const int qpk_t = ap.qpk == 1 ? 1 : (ap.qpk == 2 ? 2 : (ap.qpk <= 4 ? 4 : 8));
      ap.head_chunks = (ap.qpk + qpk_t - 1) / qpk_t;
      {  // workspace carve-up: identical to b200_attn_decode
        uint8_t* ws = reinterpret_cast<uint8_t*>(op.workspace);
        ap.ws_cnt = reinterpret_cast<unsigned int*>(ws);
        size_t o2 = (((size_t)B * op.n_head * sizeof(unsigned int)) + 255) & ~(size_t)255;
        ap.ws_ml = reinterpret_cast<float*>(ws + o2);
        o2 += (size_t)B * op.n_head * ap.nsplit * 2 * sizeof(float);
        ap.ws_acc = reinterpret_cast<float*>(ws + o2);
      }
      ap.scale_log2 = __fdiv_rn(1.0f, __fsqrt_rn((float)op.head_size)) * 1.4426950408889634f;  // == the host expression of b200_attn_decode
      const int sub = tid >> 7, stid = tid & 127;
      const int ny = op.n_groups * ap.head_chunks;
      const int items = ap.nsplit * ny * B;
      if (sub < sp.attn_nsub) {
        const uint32_t bar_id = 3 + sub;
        auto sub_sync = [bar_id]() { asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory"); };
        // item -> (CTA, sub-block): spread over CTAs first
        for (int item = blockIdx.x + gridDim.x * sub; item < items; item += gridDim.x * sp.attn_nsub) {
          const int split = item % ap.nsplit, by = (item / ap.nsplit) % ny, b = item / (ap.nsplit * ny);
          uint8_t* scratch = sb + red_a;
#define B200_ATTN_CASE(HS_, QPK_, DPL_)                                                                          \
  {                                                                                                              \
    AttnSmem<HS_, QPK_>& sm = *reinterpret_cast<AttnSmem<HS_, QPK_>*>(scratch + sub * ((sizeof(AttnSmem<HS_, QPK_>) + 15) & ~size_t(15))); \
    attn_decode_item<HS_, QPK_, DPL_>(ap, sm, stid, split, by, b, sub_sync);                                      \
  }
          if (op.head_size == 128) {
            if (qpk_t == 1) B200_ATTN_CASE(128, 1, 16)
            else if (qpk_t == 2) B200_ATTN_CASE(128, 2, 16)
            else if (qpk_t == 4) B200_ATTN_CASE(128, 4, 8)
            else B200_ATTN_CASE(128, 8, 4)
          } else {
            if (qpk_t == 1) B200_ATTN_CASE(64, 1, 16)
            else if (qpk_t == 2) B200_ATTN_CASE(64, 2, 16)
            else if (qpk_t == 4) B200_ATTN_CASE(64, 4, 8)
            else B200_ATTN_CASE(64, 8, 4)
          }
#undef B200_ATTN_CASE
          sub_sync();  // the scratch is reused by this sub-block's next item
        }
      }
    } else {
      // ============================================ EMBED op ====================================================
      if (blockIdx.x == 0) {
        const int64_t dim = op.K, vocab = op.N;
        const __nv_bfloat16* wte = reinterpret_cast<const __nv_bfloat16*>(op.w);
        __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(op.y);
        for (int b = 0; b < B; ++b) {
          int64_t tok = op.idx_is_int64 ? __ldcg(reinterpret_cast<const long long*>(op.x) + b) : (int64_t)__ldcg(reinterpret_cast<const int*>(op.x) + b);
          tok = tok < 0 ? 0 : (tok >= vocab ? vocab - 1 : tok);  // never read out of bounds
          const uint4* s4 = reinterpret_cast<const uint4*>(wte + tok * dim);
          uint4* d4 = reinterpret_cast<uint4*>(out + (int64_t)b * dim);
          for (int64_t i = tid; i < (dim >> 3); i += kDConsumers) d4[i] = s4[i];
        }
      }
    }
    stamp(oi, 2);
    if (oi + 1 < sp.n_ops) grid_sync(oi + 1, oi);
  }
}

static std::atomic<int> g_lookahead_kb{0}, g_trace{0}, g_debug{0};

template <int NB, bool INT4>
static cudaError_t launch_step(const StepParams& p, int gx, cudaStream_t st) {
  static bool configured = false;  // per instantiation
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(decode_step_kernel<NB, INT4>, cudaFuncAttributeMaxDynamicSharedMemorySize, kDDynSmem);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  // cooperative launch: the runtime guarantees (or refuses) co-residency of all CTAs, which the grid barrier needs
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)gx);
  cfg.blockDim = dim3(kDThreadsAll);
  cfg.dynamicSmemBytes = kDDynSmem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return cudaLaunchKernelEx(&cfg, decode_step_kernel<NB, INT4>, p);
}

template <int HS, int QPK>
static size_t attn_smem_bytes() { return (sizeof(AttnSmem<HS, QPK>) + 15) & ~size_t(15); }

static size_t attn_smem_for(int hs, int qpk_t) {
  if (hs == 128) return qpk_t == 1 ? attn_smem_bytes<128, 1>() : qpk_t == 2 ? attn_smem_bytes<128, 2>() : qpk_t == 4 ? attn_smem_bytes<128, 4>() : attn_smem_bytes<128, 8>();
  return qpk_t == 1 ? attn_smem_bytes<64, 1>() : qpk_t == 2 ? attn_smem_bytes<64, 2>() : qpk_t == 4 ? attn_smem_bytes<64, 4>() : attn_smem_bytes<64, 8>();
}

struct StepPlan {
  int max_k = 0, red_bufs = 1, red_cols = 8, scratch_bytes = 0, attn_nsub = 4;
};

// validates the ops and sizes the kernel's shared-memory plan; fills `plan`
static int plan_program(const b200_decode_program* prog, StepPlan* plan, const char* who) {
  B200_CHECK_ARG(prog != nullptr && prog->ops_host && prog->ops_dev && prog->sync, "%s: null program / ops / sync pointer", who);
  B200_CHECK_ARG(prog->n_ops > 0 && prog->n_ops <= 65536, "%s: bad n_ops %d", who, prog->n_ops);
  B200_CHECK_ARG(prog->sync_bytes >= 256, "%s: sync buffer must be >= 256 bytes", who);
  B200_CHECK_ARG(prog->batch >= 1 && prog->batch <= 16, "%s: batch must be in [1,16] (got %d)", who, prog->batch);
  B200_CHECK_ARG(prog->qtype == B200_Q_NF4 || prog->qtype == B200_Q_FP4 || prog->qtype == B200_Q_INT4, "%s: bad qtype %d", who, prog->qtype);
  const bool int4 = prog->qtype == B200_Q_INT4;
  B200_CHECK_ARG(prog->blocksize == (int4 ? 128 : 64), "%s: blocksize must be %d for this qtype (got %d)", who, int4 ? 128 : 64, prog->blocksize);
  const int NB = prog->batch > 8 ? 2 : 1, COLS = 8 * NB;
  size_t attn_need = 0;
  int max_k = 0, attn_sub_want = 1;
  for (int i = 0; i < prog->n_ops; ++i) {
    const b200_step_op& op = prog->ops_host[i];
    if (op.kind == B200_OP_GEMV) {
      B200_CHECK_ARG(op.w && op.x && op.y, "%s: op %d (GEMV): null w/x/y", who, i);
      B200_CHECK_ARG(op.N > 0 && op.K > 0 && op.K % (int4 ? 128 : 64) == 0 && op.K <= (1 << 20), "%s: op %d (GEMV): bad shape N=%lld K=%lld", who, i, (long long)op.N, (long long)op.K);
      B200_CHECK_ARG(op.epilogue >= B200_EPI_STORE && op.epilogue <= B200_EPI_SWIGLU, "%s: op %d: bad epilogue %d", who, i, op.epilogue);
      B200_CHECK_ARG(op.epilogue != B200_EPI_SWIGLU || op.w2, "%s: op %d: SWIGLU needs w2", who, i);
      B200_CHECK_ARG(op.epilogue != B200_EPI_RESIDUAL || op.residual, "%s: op %d: RESIDUAL epilogue needs residual", who, i);
      B200_CHECK_ARG(op.x_out == nullptr || (op.x_out != op.x && op.delta), "%s: op %d: x_out needs delta and must not alias x", who, i);
      B200_CHECK_ARG(op.y != op.x && op.y != op.residual && op.y != op.delta, "%s: op %d: y must not alias an input of the same op", who, i);
      B200_CHECK_ARG(((reinterpret_cast<uintptr_t>(op.w) | reinterpret_cast<uintptr_t>(op.w2) | reinterpret_cast<uintptr_t>(op.x) |
                       reinterpret_cast<uintptr_t>(op.delta) | reinterpret_cast<uintptr_t>(op.x_out) | reinterpret_cast<uintptr_t>(op.norm_weight)) & 15) == 0,
                     "%s: op %d: pointers must be 16-byte aligned", who, i);
      if (op.K > max_k) max_k = (int)op.K;
    } else if (op.kind == B200_OP_ATTN) {
      B200_CHECK_ARG(op.x && op.cos && op.sin && op.pos && op.k_cache && op.v_cache && op.y && op.workspace, "%s: op %d (ATTN): null pointer", who, i);
      B200_CHECK_ARG(op.n_head > 0 && op.n_groups > 0 && op.max_seq > 0 && op.n_head % op.n_groups == 0, "%s: op %d (ATTN): bad head shape", who, i);
      if (op.head_size != 64 && op.head_size != 128) {
        set_error("%s: op %d (ATTN): head_size %d not supported (64 or 128)", who, i, op.head_size);
        return B200_ERR_UNSUPPORTED;
      }
      B200_CHECK_ARG(op.rope_n_elem >= 0 && op.rope_n_elem <= op.head_size && op.rope_n_elem % 2 == 0, "%s: op %d (ATTN): bad rope_n_elem %d", who, i, op.rope_n_elem);
      B200_CHECK_ARG(op.table_dtype == B200_F32 || op.table_dtype == B200_BF16, "%s: op %d (ATTN): cos/sin must be fp32 or bf16", who, i);
      B200_CHECK_ARG(op.num_splits >= 1 && op.num_splits <= 64, "%s: op %d (ATTN): num_splits must be in [1,64] (got %d)", who, i, op.num_splits);
      const int qpk = op.n_head / op.n_groups;
      const int qpk_t = qpk == 1 ? 1 : (qpk == 2 ? 2 : (qpk <= 4 ? 4 : 8));
      const size_t need = attn_smem_for(op.head_size, qpk_t);
      if (need > attn_need) attn_need = need;
      const int items = op.num_splits * op.n_groups * ((qpk + qpk_t - 1) / qpk_t) * prog->batch;
      const int want_sub = (items + sm_count() - 1) / sm_count();
      if (want_sub > attn_sub_want) attn_sub_want = want_sub;
    } else if (op.kind == B200_OP_EMBED) {
      B200_CHECK_ARG(op.w && op.x && op.y && op.N > 0 && op.K > 0 && op.K % 8 == 0, "%s: op %d (EMBED): bad arguments", who, i);
      B200_CHECK_ARG(((reinterpret_cast<uintptr_t>(op.w) | reinterpret_cast<uintptr_t>(op.y)) & 15) == 0, "%s: op %d (EMBED): pointers must be 16-byte aligned", who, i);
    } else {
      B200_CHECK_ARG(false, "%s: op %d: unknown kind %d", who, i, op.kind);
    }
  }
  B200_CHECK_ARG(max_k > 0, "%s: a program needs at least one GEMV op", who);
  // mirror of the kernel's shared-memory carve-up (assuming the dynamic window starts within the first 1 KB)
  const size_t stage = 2 * (size_t)kDW * (int4 ? 1088 : 576);
  const int red_cols = prog->batch <= 2 ? 2 : COLS;
  const size_t red_one = (size_t)kDW * 2 * 16 * red_cols * 4;
  const int xstride = max_k + 8;
  auto ring_slots = [&](size_t scratch) -> size_t {
    size_t off = 1024 + 128 + 128 + 256;
    off += scratch + kDW * 16 * 4 + 16 * 4;
    if (int4) off += ((((size_t)prog->batch * (max_k / 64)) + 3) & ~(size_t)3) * 4;
    off += (((size_t)prog->batch * xstride * 2) + 127) & ~(size_t)127;
    const size_t end = 1024 + (size_t)kDDynSmem;
    if (int4) return off < end ? (end - off) / stage : 0;   // no table
    const size_t lut_at = (off + 0xFFFF) & ~(size_t)0xFFFF, ring_c = lut_at + 0x10000;
    if (ring_c > end) return 0;
    return (lut_at - off) / stage + (end - ring_c) / stage;
  };
  auto scratch_for = [&](int red_bufs, int nsub) -> size_t {
    size_t s = red_bufs * red_one, a = attn_need * nsub;
    return ((s > a ? s : a) + 127) & ~(size_t)127;
  };
  // prefer (2 reduction buffers, 4 attention sub-blocks); back off while the ring would get fewer than 3 stages
  const int nsub_hi = attn_sub_want > 4 ? 4 : attn_sub_want, nsub_lo = nsub_hi > 2 ? 2 : nsub_hi;
  const int cand[4][2] = {{2, nsub_hi}, {2, nsub_lo}, {1, nsub_lo}, {1, 1}};
  int pick = -1;
  for (int c = 0; c < 4; ++c) {
    if (ring_slots(scratch_for(cand[c][0], cand[c][1])) >= (c == 3 ? 2u : 3u)) { pick = c; break; }
  }
  if (pick < 0) {
    set_error("%s: batch %d x K %d does not fit the kernel's shared memory plan", who, prog->batch, max_k);
    return B200_ERR_UNSUPPORTED;
  }
  plan->max_k = max_k;
  plan->red_bufs = cand[pick][0];
  plan->red_cols = red_cols;
  plan->attn_nsub = cand[pick][1];
  plan->scratch_bytes = (int)scratch_for(cand[pick][0], cand[pick][1]);
  return B200_OK;
}

}  // namespace b200

using namespace b200;

extern "C" {

int b200_decode_step_tune(int lookahead_kb, int trace) {
  g_lookahead_kb = lookahead_kb < 0 ? 0 : lookahead_kb;
  g_trace = trace & 1;
  g_debug = trace >> 4;   // experiments: trace bits 4.. = kernel debug bits (skip compute, split copies)
  return B200_OK;
}

int b200_decode_program_upload(const b200_decode_program* prog, b200_stream_t stream) {
  StepPlan plan;
  int rc = plan_program(prog, &plan, "b200_decode_program_upload");
  if (rc != B200_OK) return rc;
  B200_CUDA(cudaMemcpyAsync(const_cast<b200_step_op*>(prog->ops_dev), prog->ops_host, sizeof(b200_step_op) * (size_t)prog->n_ops,
                            cudaMemcpyHostToDevice, (cudaStream_t)stream));
  return B200_OK;
}

int b200_decode_step(const b200_decode_program* prog, b200_stream_t stream) {
  StepPlan plan;
  int rc = plan_program(prog, &plan, "b200_decode_step");
  if (rc != B200_OK) return rc;
  StepParams p{};
  p.ops = prog->ops_dev;
  p.n_ops = prog->n_ops;
  p.batch = prog->batch;
  p.qtype = prog->qtype;
  p.max_k = plan.max_k;
  p.xstride = plan.max_k + 8;
  p.red_bufs = plan.red_bufs;
  p.red_cols = plan.red_cols;
  p.scratch_bytes = plan.scratch_bytes;
  p.attn_nsub = plan.attn_nsub;
  p.sync = reinterpret_cast<unsigned int*>(prog->sync);
  p.lookahead_kb = g_lookahead_kb.load();
  p.debug = g_debug.load();
  p.trace = (g_trace.load() && prog->sync_bytes >= 256 + 64 * (int64_t)prog->n_ops) ? reinterpret_cast<long long*>(reinterpret_cast<uint8_t*>(prog->sync) + 256) : nullptr;
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA(cudaMemsetAsync(prog->sync, 0, 8, st));
  const int gx = sm_count();
  const bool int4 = prog->qtype == B200_Q_INT4;
  const int NB = prog->batch > 8 ? 2 : 1;
  cudaError_t e;
  if (int4) e = NB == 2 ? launch_step<2, true>(p, gx, st) : launch_step<1, true>(p, gx, st);
  else e = NB == 2 ? launch_step<2, false>(p, gx, st) : launch_step<1, false>(p, gx, st);
  if (e != cudaSuccess) {
    set_error("b200_decode_step: launch failed: %s", cudaGetErrorString(e));
    return B200_ERR_CUDA;
  }
  return B200_OK;
}

}  // extern "C"
