// Body of the fused RoPE + KV-append + split-K decode attention for ONE work item (batch b, KV group / head chunk,
// split), written for a group of 128 threads.  Shared by the stand-alone kernel (attn_decode.cu: one CTA per
// item, __syncthreads) and by the persistent decode-step kernel (decode_step.cu: up to four 128-thread
// sub-blocks of a CTA, each on its own named barrier).  See attn_decode.cu for the reference lines it replaces.
#pragma once
#include "common.cuh"

namespace b200 {

struct AttnParams {
  const __nv_bfloat16* qkv;
  const void* cos;
  const void* sin;
  int table_bf16;
  int table_presel;
  int rope_n;
  const void* pos;
  int pos_i64;
  __nv_bfloat16* kc;
  __nv_bfloat16* vc;
  __nv_bfloat16* out;
  int batch, n_head, n_groups, hs, max_seq, qpk;
  int nsplit;
  int head_chunks;  // ceil(qpk / QPK)
  float* ws_acc;    // [batch][n_head][nsplit][hs]
  float* ws_ml;     // [batch][n_head][nsplit][2]
  unsigned int* ws_cnt;  // [batch][n_groups * head_chunks]
  float scale_log2;
  unsigned int* fuse_cnt;   // fused into the QKV GEMV (qgemv_v4.cu): per-head row-tile counters, cleared by the merging sub-block
};

// validates `a` for the decode attention fused into the QKV GEMV (qgemv_v4.cu: batch 1, n_head == n_groups, head size 128,
// splits as sub-blocks of one CTA) and fills the kernel parameters; B200_ERR_UNSUPPORTED (no message) when it does not apply
int attn_fused_params(const b200_attn_decode_args* a, int nsplit, AttnParams* out);
// grouped-query kernel on mma.sync (attn_decode_gqa.cu)
bool attn_decode_gqa_supported(int head_size, int qpk);
cudaError_t launch_attn_decode_gqa(const AttnParams& p, dim3 grid, cudaStream_t st);

__device__ __forceinline__ float table_at(const void* tab, int is_bf16, int64_t i) {
  return is_bf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(tab)[i]) : reinterpret_cast<const float*>(tab)[i];
}

// apply_rope (lit_gpt/model.py:767-773) on one head vector held in shared memory as bf16-exact floats.
// dst[d] = bf16(x[d]*cos[d] + rot[d]*sin[d]) for d < n, x[d] otherwise; rot = cat(-x2, x1).
__device__ __forceinline__ float rope_elem(const float* x, int d, int n, const void* cos, const void* sin, int tb, int64_t row) {
  if (d >= n) return x[d];
  const int half = n >> 1;
  const float c = table_at(cos, tb, row * n + d), s = table_at(sin, tb, row * n + d);
  const float rot = d < half ? -x[d + half] : x[d - half];
  if (tb) return round_bf16(__fadd_rn(round_bf16(__fmul_rn(x[d], c)), round_bf16(__fmul_rn(rot, s))));  // bf16 tables: 3 roundings
  return round_bf16(__fadd_rn(__fmul_rn(x[d], c), __fmul_rn(rot, s)));                                     // fp32 tables: 1 rounding
}

template <int N>
__device__ __forceinline__ void load_row(uint32_t (&r)[N], const __nv_bfloat16* p) {
  if constexpr (N == 8) {
    uint4 a = *reinterpret_cast<const uint4*>(p), b = *reinterpret_cast<const uint4*>(p + 8);
    r[0] = a.x; r[1] = a.y; r[2] = a.z; r[3] = a.w; r[4] = b.x; r[5] = b.y; r[6] = b.z; r[7] = b.w;
  } else if constexpr (N == 4) {
    uint4 a = *reinterpret_cast<const uint4*>(p);
    r[0] = a.x; r[1] = a.y; r[2] = a.z; r[3] = a.w;
  } else {
    uint2 a = *reinterpret_cast<const uint2*>(p);
    r[0] = a.x; r[1] = a.y;
  }
}

// HS head size, QPK query heads handled per CTA (shared K/V rows), DPL dims per lane (HS/DPL lanes per key)
template <int HS, int QPK>
struct AttnSmem {
  float q[QPK][HS];
  float k[HS], v[HS], tmp[HS];
  float acc[4][QPK][HS];
  float m[4][QPK], l[4][QPK];
  float pm[QPK], pl[QPK];   // this CTA's split partial (m, l) when the splits merge through distributed shared memory
  unsigned int last;
};

// activations written earlier by OTHER CTAs of the same (persistent) kernel must not be served from a stale L1 line
__device__ __forceinline__ float ld_act_bf16(const __nv_bfloat16* p) {
  return __bfloat162float(__ushort_as_bfloat16(__ldcg(reinterpret_cast<const unsigned short*>(p))));
}

// tid in [0, 128); `sync()` is a barrier over exactly those 128 threads; `by` = group * head_chunks + head chunk
// CLUSTER: the nsplit CTAs of one (batch, group, head chunk) form a thread-block cluster (launch attribute); their split
// partials stay in shared memory and cluster rank 0 merges them through distributed shared memory -- no workspace, no
// fence, no ticket, three dependent L2 round trips fewer than the workspace protocol.
// INTRA: the nsplit (<= 4) splits of one item are 128-thread sub-blocks of ONE CTA (AttnSmem instances contiguous in shared
// memory, `sm` being this sub-block's); `syncall()` is a barrier over all of them; sub-block 0 merges from shared memory.
struct NoSync {
  __device__ __forceinline__ void operator()() const {}
};
template <int HS, int QPK, int DPL, typename Sync, bool CLUSTER = false, bool INTRA = false, typename SyncAll = NoSync, int KPI_ = 4>
__device__ __forceinline__ void attn_decode_item(const AttnParams& p, AttnSmem<HS, QPK>& sm, const int tid, const int split, const int by,
                                                 const int b, Sync sync, SyncAll syncall = SyncAll{}) {
  constexpr int LPK = HS / DPL, GPW = 32 / LPK, SLOTS = 4 * GPW, NR = DPL / 2;
  float (&s_q)[QPK][HS] = sm.q;
  float (&s_k)[HS] = sm.k;
  float (&s_v)[HS] = sm.v;
  float (&s_tmp)[HS] = sm.tmp;
  float (&s_acc)[4][QPK][HS] = sm.acc;
  float (&s_m)[4][QPK] = sm.m;
  float (&s_l)[4][QPK] = sm.l;
  unsigned int& s_last = sm.last;

  const int lane = tid & 31, warp = tid >> 5;
  const int gi = by / p.head_chunks, hc = by % p.head_chunks;
  const int h0 = hc * QPK;                                     // first query head (within the group) of this item
  const int nh = min(QPK, p.qpk - h0);                         // live heads

  const bool last_split = split == p.nsplit - 1;

  // ---- q (and, in the last split, the new k / v): load, RoPE, keep as bf16-exact floats in smem -----------
  // All global loads of this phase are issued together (q heads, k, v next to the position load above) and the
  // cos / sin entries are fetched once per thread, so the phase costs two dependent memory round trips and two
  // barriers regardless of the number of heads.  The raw vectors are staged in s_acc, unused until the final merge.
  const __nv_bfloat16* grp = p.qkv + ((int64_t)b * p.n_groups + gi) * (p.qpk + 2) * HS;
  float (&raw_q)[QPK][HS] = s_acc[0];
  float* raw_k = s_acc[1][0];
  float rq[QPK], rk = 0.f, rv = 0.f;
  if (tid < HS) {
#pragma unroll
    for (int h = 0; h < QPK; ++h) rq[h] = h < nh ? ld_act_bf16(grp + (h0 + h) * HS + tid) : 0.f;
    if (last_split) { rk = ld_act_bf16(grp + p.qpk * HS + tid); rv = ld_act_bf16(grp + (p.qpk + 1) * HS + tid); }
  }
  // the position is loaded AFTER the activation loads were issued (they do not depend on it): one round trip, not two
  int64_t pos = p.pos_i64 ? __ldcg(reinterpret_cast<const long long*>(p.pos)) : (int64_t)__ldcg(reinterpret_cast<const int*>(p.pos));
  pos = pos < 0 ? 0 : (pos >= p.max_seq ? p.max_seq - 1 : pos);  // host validates; never index out of bounds
  if (tid < HS) {
#pragma unroll
    for (int h = 0; h < QPK; ++h) raw_q[h][tid] = rq[h];
    raw_k[tid] = rk;
    s_v[tid] = rv;
  }
  sync();
  if (tid < HS) {
    const int d = tid, n = p.rope_n, half = n >> 1;
    float c = 1.f, sn = 0.f;
    if (d < n) {
      const int64_t row = p.table_presel ? 0 : pos;
      c = table_at(p.cos, p.table_bf16, row * n + d);
      sn = table_at(p.sin, p.table_bf16, row * n + d);
    }
    // apply_rope (lit_gpt/model.py:767-773), same rounding points as rope_elem()
    auto rope = [&](const float* x) -> float {
      if (d >= n) return x[d];
      const float rot = d < half ? -x[d + half] : x[d - half];
      if (p.table_bf16) return round_bf16(__fadd_rn(round_bf16(__fmul_rn(x[d], c)), round_bf16(__fmul_rn(rot, sn))));
      return round_bf16(__fadd_rn(__fmul_rn(x[d], c), __fmul_rn(rot, sn)));
    };
#pragma unroll
    for (int h = 0; h < QPK; ++h)
      if (h < nh) s_q[h][d] = rope(raw_q[h]);
    if (last_split) {
      const float kd = rope(raw_k);
      s_k[d] = kd;
      if (hc == 0) {  // exactly one CTA per (batch, group) appends to the cache (KVCache.forward, model.py:793-794)
        const int64_t off = (((int64_t)b * p.n_groups + gi) * p.max_seq + pos) * HS;
        p.kc[off + d] = __float2bfloat16_rn(kd);
        p.vc[off + d] = __float2bfloat16_rn(s_v[d]);
      }
    }
  }
  sync();

  const int gidx = lane / LPK, dl = lane % LPK;
  float q[QPK][DPL];
#pragma unroll
  for (int h = 0; h < QPK; ++h)
#pragma unroll
    for (int i = 0; i < DPL; ++i) q[h][i] = h < nh ? s_q[h][dl * DPL + i] * p.scale_log2 : 0.f;

  float m[QPK], l[QPK], acc[QPK][DPL];
#pragma unroll
  for (int h = 0; h < QPK; ++h) {
    m[h] = -INFINITY; l[h] = 0.f;
#pragma unroll
    for (int i = 0; i < DPL; ++i) acc[h][i] = 0.f;
  }

  auto update = [&](const uint32_t (&kr)[NR], const uint32_t (&vr)[NR], bool valid) {
    float kf[DPL], vf[DPL];
#pragma unroll
    for (int i = 0; i < NR; ++i) { kf[2 * i] = bf16lo(kr[i]); kf[2 * i + 1] = bf16hi(kr[i]); vf[2 * i] = bf16lo(vr[i]); vf[2 * i + 1] = bf16hi(vr[i]); }
#pragma unroll
    for (int h = 0; h < QPK; ++h) {
      float sc = 0.f;
#pragma unroll
      for (int i = 0; i < DPL; ++i) sc = fmaf(q[h][i], kf[i], sc);
#pragma unroll
      for (int o = LPK / 2; o > 0; o >>= 1) sc += __shfl_xor_sync(0xffffffffu, sc, o);
      if (valid) {  // uniform within the LPK lanes of a key
        const float mn = fmaxf(m[h], sc);
        const float corr = exp2f(m[h] - mn);  // m = -inf first time: exp2(-inf) = 0
        const float pr = exp2f(sc - mn);
        l[h] = fmaf(l[h], corr, pr);
#pragma unroll
        for (int i = 0; i < DPL; ++i) acc[h][i] = fmaf(acc[h][i], corr, pr * vf[i]);
        m[h] = mn;
      }
    }
  };

  // ---- cached keys [kbeg, kend) of this split -----------------------------------------------------------
  const int64_t per = (pos + p.nsplit - 1) / p.nsplit;
  const int64_t kbeg = (int64_t)split * per, kend = min(pos, kbeg + per);
  const __nv_bfloat16* kbase = p.kc + ((int64_t)b * p.n_groups + gi) * p.max_seq * HS + dl * DPL;
  const __nv_bfloat16* vbase = p.vc + ((int64_t)b * p.n_groups + gi) * p.max_seq * HS + dl * DPL;
  constexpr int KPI = KPI_;  // keys per lane group and iteration: all their K / V rows are in flight before the first update
  for (int64_t base = kbeg + warp * GPW; base < kend; base += KPI * SLOTS) {  // warp-uniform trip count
    uint32_t kr[KPI][NR], vr[KPI][NR];
    bool ok[KPI];
#pragma unroll
    for (int u = 0; u < KPI; ++u) {
      const int64_t kk = base + gidx + (int64_t)u * SLOTS;
      ok[u] = kk < kend;
#pragma unroll
      for (int i = 0; i < NR; ++i) kr[u][i] = vr[u][i] = 0u;
      if (ok[u]) { load_row<NR>(kr[u], kbase + kk * HS); load_row<NR>(vr[u], vbase + kk * HS); }
    }
#pragma unroll
    for (int u = 0; u < KPI; ++u) {
      if (base + (int64_t)u * SLOTS < kend) update(kr[u], vr[u], ok[u]);  // warp-uniform condition (shuffles inside)
    }
  }
  // ---- the key being appended comes from shared memory (never re-read from HBM) ---------------------------
  if (last_split && warp == 0) {
    uint32_t kr[NR], vr[NR];
#pragma unroll
    for (int i = 0; i < NR; ++i) {
      kr[i] = pack_bf16(s_k[dl * DPL + 2 * i], s_k[dl * DPL + 2 * i + 1]);
      vr[i] = pack_bf16(s_v[dl * DPL + 2 * i], s_v[dl * DPL + 2 * i + 1]);
    }
    update(kr, vr, gidx == 0);
  }

  // ---- merge lane groups (shuffles), then warps (smem) ----------------------------------------------------
#pragma unroll
  for (int h = 0; h < QPK; ++h) {
#pragma unroll
    for (int o = LPK; o < 32; o <<= 1) {
      const float mo = __shfl_xor_sync(0xffffffffu, m[h], o), lo = __shfl_xor_sync(0xffffffffu, l[h], o);
      const float mn = fmaxf(m[h], mo);
      const float ca = mn == -INFINITY ? 0.f : exp2f(m[h] - mn), cb = mn == -INFINITY ? 0.f : exp2f(mo - mn);
      l[h] = l[h] * ca + lo * cb;
#pragma unroll
      for (int i = 0; i < DPL; ++i) acc[h][i] = acc[h][i] * ca + __shfl_xor_sync(0xffffffffu, acc[h][i], o) * cb;
      m[h] = mn;
    }
    if (gidx == 0) {
#pragma unroll
      for (int i = 0; i < DPL; ++i) s_acc[warp][h][dl * DPL + i] = acc[h][i];
      if (dl == 0) { s_m[warp][h] = m[h]; s_l[warp][h] = l[h]; }
    }
  }
  sync();

  // each thread finalises dims d = tid (+128) for every live head: merge the 4 warps
  const int H = p.n_head;
  for (int h = 0; h < nh; ++h) {
    const int head = gi * p.qpk + h0 + h;
    float mn = fmaxf(fmaxf(s_m[0][h], s_m[1][h]), fmaxf(s_m[2][h], s_m[3][h]));
    float c[4], lt = 0.f;
#pragma unroll
    for (int w = 0; w < 4; ++w) { c[w] = mn == -INFINITY ? 0.f : exp2f(s_m[w][h] - mn); lt += s_l[w][h] * c[w]; }
    for (int d = tid; d < HS; d += 128) {
      float a = 0.f;
#pragma unroll
      for (int w = 0; w < 4; ++w) a += s_acc[w][h][d] * c[w];
      if (p.nsplit == 1) p.out[((int64_t)b * H + head) * HS + d] = __float2bfloat16_rn(a / lt);
      else if (CLUSTER || INTRA) s_q[h][d] = a;   // q lives in registers by now: its shared-memory rows carry the split partial
      else __stcg(p.ws_acc + (((int64_t)b * H + head) * p.nsplit + split) * HS + d, a);
    }
    if (p.nsplit > 1 && tid == 0) {
      if (CLUSTER || INTRA) { sm.pm[h] = mn; sm.pl[h] = lt; }
      else {
        float* ml = p.ws_ml + (((int64_t)b * H + head) * p.nsplit + split) * 2;
        __stcg(ml, mn);
        __stcg(ml + 1, lt);
      }
    }
  }
  if (p.nsplit == 1) {
    if (p.fuse_cnt && tid == 0) p.fuse_cnt[by] = 0u;
    return;
  }
  if constexpr (INTRA) {
    syncall();   // every sub-block's partial (m, l, o) is in shared memory
    if (split == 0) {
      const int NS = p.nsplit;
      const AttnSmem<HS, QPK>* sib = &sm;   // sub-block r of this item: sib[r]
      for (int i = tid; i < nh * HS; i += 128) {
        const int h = i / HS, d = i % HS;
        float mn = -INFINITY, lt = 0.f, a = 0.f;
        for (int r = 0; r < NS; ++r) mn = fmaxf(mn, sib[r].pm[h]);
        for (int r = 0; r < NS; ++r) lt += sib[r].pm[h] == -INFINITY ? 0.f : sib[r].pl[h] * exp2f(sib[r].pm[h] - mn);
        const float inv = 1.f / lt;
        for (int r = 0; r < NS; ++r) a = fmaf(sib[r].q[h][d], sib[r].pm[h] == -INFINITY ? 0.f : exp2f(sib[r].pm[h] - mn) * inv, a);
        p.out[((int64_t)b * H + gi * p.qpk + h0 + h) * HS + d] = __float2bfloat16_rn(a);
      }
      if (p.fuse_cnt && tid == 0) p.fuse_cnt[by] = 0u;   // every sub-block is past its wait on the counter
    }
    return;
  }
  if constexpr (CLUSTER) {
    // cluster barrier #1: every split's partial is in its CTA's shared memory
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (split == 0) {
      const int NS = p.nsplit;
      const uint32_t q_a = (uint32_t)__cvta_generic_to_shared(&s_q[0][0]);
      const uint32_t pm_a = (uint32_t)__cvta_generic_to_shared(&sm.pm[0]), pl_a = (uint32_t)__cvta_generic_to_shared(&sm.pl[0]);
      auto remote = [](uint32_t addr, int rank) -> float {
        uint32_t ra;
        float v;
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(addr), "r"(rank));
        asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(ra) : "memory");
        return v;
      };
      for (int i = tid; i < nh * HS; i += 128) {
        const int h = i / HS, d = i % HS;
        float ms[8], ls[8], vs[8];
#pragma unroll
        for (int r = 0; r < 8; ++r) {
          ms[r] = r < NS ? remote(pm_a + h * 4, r) : -INFINITY;
          ls[r] = r < NS ? remote(pl_a + h * 4, r) : 0.f;
          vs[r] = r < NS ? remote(q_a + (h * HS + d) * 4, r) : 0.f;
        }
        float mn = -INFINITY, lt = 0.f, a = 0.f;
#pragma unroll
        for (int r = 0; r < 8; ++r) mn = fmaxf(mn, ms[r]);
#pragma unroll
        for (int r = 0; r < 8; ++r) lt += ms[r] == -INFINITY ? 0.f : ls[r] * exp2f(ms[r] - mn);
        const float inv = 1.f / lt;
#pragma unroll
        for (int r = 0; r < 8; ++r) a = fmaf(vs[r], ms[r] == -INFINITY ? 0.f : exp2f(ms[r] - mn) * inv, a);
        p.out[((int64_t)b * H + gi * p.qpk + h0 + h) * HS + d] = __float2bfloat16_rn(a);
      }
    }
    // cluster barrier #2: no CTA leaves (and frees its shared memory) before rank 0 has read it
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    return;
  }

  // ---- last CTA of this (batch, group, head chunk) merges the splits ---------------------------------------
  __threadfence();
  sync();
  if (tid == 0) {
    unsigned int* cnt = p.ws_cnt + (int64_t)b * p.n_groups * p.head_chunks + by;
    const unsigned int old = atomicAdd(cnt, 1u);
    s_last = old == (unsigned int)(p.nsplit - 1);
    if (s_last) { *cnt = 0u; if (p.fuse_cnt) p.fuse_cnt[by] = 0u; }  // self-cleaning
  }
  sync();
  if (!s_last) return;
  __threadfence();
  // (m, l) of every (head, split) staged in shared memory (s_acc is free again), converted in place to the split's
  // weight exp2(m_s - m) / l_total; then one (head, dim) pair per thread with the split loads four deep in flight
  float* s_w = &s_acc[0][0][0];                  // [nh][nsplit] weights   (nsplit <= 64: 8 * 64 * 2 floats fit easily)
  float* s_mraw = s_w + QPK * 64;
  const int NS = p.nsplit;
  for (int i = tid; i < nh * NS; i += 128) {
    const int h = i / NS, sp = i % NS;
    const float* ml = p.ws_ml + (((int64_t)b * H + gi * p.qpk + h0 + h) * NS + sp) * 2;
    s_mraw[i] = __ldcg(ml);
    s_w[i] = __ldcg(ml + 1);
  }
  sync();
  if (tid < nh) {
    float mn = -INFINITY, lt = 0.f;
    for (int sp = 0; sp < NS; ++sp) mn = fmaxf(mn, s_mraw[tid * NS + sp]);
    for (int sp = 0; sp < NS; ++sp) lt += s_w[tid * NS + sp] * exp2f(s_mraw[tid * NS + sp] - mn);
    const float inv = 1.f / lt;
    for (int sp = 0; sp < NS; ++sp) s_w[tid * NS + sp] = exp2f(s_mraw[tid * NS + sp] - mn) * inv;
  }
  sync();
  for (int i = tid; i < nh * HS; i += 128) {
    const int h = i / HS, d = i % HS;
    const int head = gi * p.qpk + h0 + h;
    const float* src = p.ws_acc + ((int64_t)b * H + head) * NS * HS + d;
    const float* w = s_w + h * NS;
    float a = 0.f;
    int sp = 0;
    for (; sp + 4 <= NS; sp += 4) {
      const float v0 = __ldcg(src + (int64_t)sp * HS), v1 = __ldcg(src + (int64_t)(sp + 1) * HS);
      const float v2 = __ldcg(src + (int64_t)(sp + 2) * HS), v3 = __ldcg(src + (int64_t)(sp + 3) * HS);
      a = fmaf(v0, w[sp], a); a = fmaf(v1, w[sp + 1], a); a = fmaf(v2, w[sp + 2], a); a = fmaf(v3, w[sp + 3], a);
    }
    for (; sp < NS; ++sp) a = fmaf(__ldcg(src + (int64_t)sp * HS), w[sp], a);
    p.out[((int64_t)b * H + head) * HS + d] = __float2bfloat16_rn(a);
  }
}

}  // namespace b200
