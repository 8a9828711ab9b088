// Grouped-query / multi-query decode attention (T = 1) on mma.sync: the q_per_kv query heads of one KV group are the
// M rows of the tensor-core tile, so each K / V row is read once and serves every head of its group.
//
// Same contract as attn_decode_kernel (attn_decode.cu; replaces lit_gpt/model.py:624-653 for one decode step): RoPE
// of q and of the new k from the reference's cos / sin tables with apply_rope's rounding points (model.py:767-773),
// append of the new k / v row to the GQA-native cache (KVCache.forward, model.py:793-794), softmax(q k^T / sqrt(hs))
// over cache positions 0..pos only (the reference masks the rest, model.py:509), fp32 running-maximum softmax, split
// over the keys across CTAs and merged by the last CTA of a (batch, group) to arrive (same workspace protocol).
// Used when q_per_kv >= 2 and head_size = 128 (Mistral-7B: 4, Llama-2-70B: 8, also per rank under tensor
// parallelism); the scalar kernel keeps multi-head attention (q_per_kv = 1) and head_size 64.
//
// One CTA = 4 warps; warp w takes the 32-key tiles w, w+4, .. of the split's range, double-buffered in its private
// shared-memory stages by cp.async; S = Q K^T and O += P V on mma.sync.m16n8k16 with ldmatrix operand loads from
// XOR-swizzled tiles (mma_tile.cuh); the new key is patched into its tile from shared memory, never re-read from HBM.
#include "attn_body.cuh"
#include "mma_tile.cuh"

namespace b200 {

namespace {

constexpr int kGqHS = 128;
constexpr int kGqRows = 16;                 // MMA M: query heads of the group (rows >= q_per_kv are zero)
constexpr int kGqTile = 32;                 // keys per warp tile
constexpr int kGqStage = 2 * kGqTile * kGqHS * 2;         // K + V tile, bytes
constexpr int kGqWarpBytes = 2 * kGqStage;                // two stages per warp (32 KB)
constexpr int kGqQBytes = kGqRows * kGqHS * 2;
constexpr int kGqSmem = 4 * kGqWarpBytes + kGqQBytes + 2 * kGqHS * 2;

// CLUSTER: the nsplit (<= 8) CTAs of a (batch, group) are one thread-block cluster; split partials stay in shared memory
// and cluster rank 0 merges them through distributed shared memory (no workspace, fence or ticket).
template <bool CLUSTER>
__global__ void __launch_bounds__(128, 1) attn_decode_gqa_kernel(const AttnParams p) {
  extern __shared__ __align__(128) uint8_t gq_smem[];
  constexpr int HS = kGqHS;
  __shared__ float s_m[4][kGqRows], s_l[4][kGqRows];
  __shared__ float s_pm[kGqRows], s_pl[kGqRows];
  __shared__ unsigned int s_last;
  uint8_t* const sQ = gq_smem + 4 * kGqWarpBytes;                          // [16][HS] bf16, swizzled
  __nv_bfloat16* const s_nk = reinterpret_cast<__nv_bfloat16*>(sQ + kGqQBytes);  // new k row (roped), new v row
  __nv_bfloat16* const s_nv = s_nk + HS;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, tq = lane & 3;
  const int split = blockIdx.x, gi = blockIdx.y, b = blockIdx.z;
  const int nh = p.qpk;
  griddep_launch();
  griddep_wait();

  // ---- q heads (+ the new k / v in the owner): load, RoPE, bf16 rows in shared memory ------------------------
  // the q loads do not depend on the position: they are issued first so that they overlap the position round trip
  const __nv_bfloat16* grp = p.qkv + ((int64_t)b * p.n_groups + gi) * (p.qpk + 2) * HS;
  const int d = tid, n = p.rope_n, half = n >> 1;
  const int dp = d < n ? (d < half ? d + half : d - half) : d;      // rotation partner
  float xq[kGqRows], xp[kGqRows], kx = 0.f, kp = 0.f, vx = 0.f;
#pragma unroll
  for (int h = 0; h < kGqRows; ++h) {
    xq[h] = h < nh ? ld_act_bf16(grp + h * HS + d) : 0.f;
    xp[h] = h < nh ? ld_act_bf16(grp + h * HS + dp) : 0.f;
  }
  kx = ld_act_bf16(grp + nh * HS + d); kp = ld_act_bf16(grp + nh * HS + dp); vx = ld_act_bf16(grp + (nh + 1) * HS + d);
  int64_t pos = p.pos_i64 ? __ldcg(reinterpret_cast<const long long*>(p.pos)) : (int64_t)__ldcg(reinterpret_cast<const int*>(p.pos));
  pos = pos < 0 ? 0 : (pos >= p.max_seq ? p.max_seq - 1 : pos);
  // keys 0..pos (the new one included) are cut into nsplit contiguous ranges; the split that owns `pos` appends it
  const int nkeys = (int)pos + 1;
  const int per = (nkeys + p.nsplit - 1) / p.nsplit;
  const int kbeg = min(split * per, nkeys), kend = min(nkeys, kbeg + per);
  const bool owner = split == (int)(pos / per);
  // the first key tile only needs the position: its copies fly while q is roped
  const __nv_bfloat16* kg = p.kc + ((int64_t)b * p.n_groups + gi) * (int64_t)p.max_seq * HS;
  const __nv_bfloat16* vg = p.vc + ((int64_t)b * p.n_groups + gi) * (int64_t)p.max_seq * HS;
  const uint32_t wbase = pf_smem(gq_smem) + warp * kGqWarpBytes;
  const int ntiles = (kend - kbeg + kGqTile - 1) / kGqTile;

  auto load_tile = [&](int tile, int buf) {   // 32 keys x 256 B of K and of V, one 256 B row per 16 lanes
    const int j0 = kbeg + tile * kGqTile;
    const uint32_t kb = wbase + buf * kGqStage, vb = kb + kGqTile * HS * 2;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int r = 2 * j + (lane >> 4), c = lane & 15;
      const bool ok = j0 + r < kend;
      const int64_t off = (int64_t)(ok ? j0 + r : 0) * HS + c * 8;
      pf_cp16(kb + sw<HS>(r, c), kg + off, ok);
      pf_cp16(vb + sw<HS>(r, c), vg + off, ok);
    }
    pf_commit();
  };

  if (warp < ntiles) load_tile(warp, 0);
  {
    float c = 1.f, sn = 0.f;
    if (d < n) {
      const int64_t row = p.table_presel ? 0 : pos;
      c = table_at(p.cos, p.table_bf16, row * n + d);
      sn = table_at(p.sin, p.table_bf16, row * n + d);
    }
    auto rope = [&](float x, float partner) -> float {   // apply_rope (model.py:767-773), rounding points of rope_elem()
      if (d >= n) return x;
      const float rot = d < half ? -partner : partner;
      if (p.table_bf16) return round_bf16(__fadd_rn(round_bf16(__fmul_rn(x, c)), round_bf16(__fmul_rn(rot, sn))));
      return round_bf16(__fadd_rn(__fmul_rn(x, c), __fmul_rn(rot, sn)));
    };
#pragma unroll
    for (int h = 0; h < kGqRows; ++h) {
      const float v = h < nh ? rope(xq[h], xp[h]) : 0.f;
      *reinterpret_cast<__nv_bfloat16*>(sQ + sw<HS>(h, d >> 3) + (d & 7) * 2) = __float2bfloat16_rn(v);
    }
    if (owner) {
      const __nv_bfloat16 kd = __float2bfloat16_rn(rope(kx, kp)), vd = __float2bfloat16_rn(vx);
      s_nk[d] = kd;
      s_nv[d] = vd;
      const int64_t off = (((int64_t)b * p.n_groups + gi) * p.max_seq + pos) * HS;
      p.kc[off + d] = kd;
      p.vc[off + d] = vd;
    }
  }
  __syncthreads();

  uint32_t qf[HS / 16][4];
#pragma unroll
  for (int ks = 0; ks < HS / 16; ++ks) ldsm_x4(qf[ks], pf_smem(sQ) + sw<HS>(lane & 15, ks * 2 + (lane >> 4)));

  float o_acc[HS / 8][4];
#pragma unroll
  for (int i = 0; i < HS / 8; ++i) o_acc[i][0] = o_acc[i][1] = o_acc[i][2] = o_acc[i][3] = 0.f;
  float mA = -INFINITY, mB = -INFINITY, lA = 0.f, lB = 0.f;   // m in the scaled (log2) domain

  int it = 0;
  for (int tile = warp; tile < ntiles; tile += 4, ++it) {
    const int buf = it & 1;
    if (tile + 4 < ntiles) { load_tile(tile + 4, buf ^ 1); pf_wait<1>(); } else { pf_wait<0>(); }
    __syncwarp();
    const int j0 = kbeg + tile * kGqTile;
    const uint32_t kb = wbase + buf * kGqStage, vb = kb + kGqTile * HS * 2;
    if (owner && (int)pos >= j0 && (int)pos < j0 + kGqTile) {   // the row being appended comes from shared memory
      const int r = (int)pos - j0, c = lane & 15;
      const uint4 val = *reinterpret_cast<const uint4*>((lane < 16 ? s_nk : s_nv) + c * 8);
      uint32_t dst = (lane < 16 ? kb : vb) + sw<HS>(r, c);
      asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(dst), "r"(val.x), "r"(val.y), "r"(val.z), "r"(val.w) : "memory");
      __syncwarp();
    }
    float s[kGqTile / 8][4];
#pragma unroll
    for (int i = 0; i < kGqTile / 8; ++i) s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f;
#pragma unroll
    for (int ks = 0; ks < HS / 16; ++ks) {
#pragma unroll
      for (int np = 0; np < kGqTile / 16; ++np) {
        uint32_t kf[4];
        ldsm_x4(kf, kb + sw<HS>(np * 16 + ((lane >> 4) << 3) + (lane & 7), ks * 2 + ((lane >> 3) & 1)));
        pf_mma(s[2 * np], qf[ks], kf[0], kf[1]);
        pf_mma(s[2 * np + 1], qf[ks], kf[2], kf[3]);
      }
    }
    float tmA = -INFINITY, tmB = -INFINITY;
#pragma unroll
    for (int i = 0; i < kGqTile / 8; ++i) {
      const int j = j0 + i * 8 + 2 * tq;
#pragma unroll
      for (int e = 0; e < 4; ++e) s[i][e] = (j + (e & 1) < kend) ? s[i][e] * p.scale_log2 : -INFINITY;
      tmA = fmaxf(tmA, fmaxf(s[i][0], s[i][1]));
      tmB = fmaxf(tmB, fmaxf(s[i][2], s[i][3]));
    }
    tmA = fmaxf(tmA, __shfl_xor_sync(0xffffffffu, tmA, 1)); tmA = fmaxf(tmA, __shfl_xor_sync(0xffffffffu, tmA, 2));
    tmB = fmaxf(tmB, __shfl_xor_sync(0xffffffffu, tmB, 1)); tmB = fmaxf(tmB, __shfl_xor_sync(0xffffffffu, tmB, 2));
    const float nmA = fmaxf(mA, tmA), nmB = fmaxf(mB, tmB);     // finite: every tile holds at least one valid key
    const float cA = exp2f(mA - nmA), cB = exp2f(mB - nmB);     // exp2(-inf) = 0 the first time
    mA = nmA; mB = nmB;
    float rsA = 0.f, rsB = 0.f;
    uint32_t pfr[kGqTile / 16][4];
#pragma unroll
    for (int i = 0; i < kGqTile / 8; ++i) {
      const float p0 = exp2f(s[i][0] - nmA), p1 = exp2f(s[i][1] - nmA), p2 = exp2f(s[i][2] - nmB), p3 = exp2f(s[i][3] - nmB);
      rsA += p0 + p1; rsB += p2 + p3;
      pfr[i >> 1][(i & 1) * 2 + 0] = pack_bf16(p0, p1);
      pfr[i >> 1][(i & 1) * 2 + 1] = pack_bf16(p2, p3);
    }
    lA = lA * cA + rsA; lB = lB * cB + rsB;
#pragma unroll
    for (int i = 0; i < HS / 8; ++i) { o_acc[i][0] *= cA; o_acc[i][1] *= cA; o_acc[i][2] *= cB; o_acc[i][3] *= cB; }
#pragma unroll
    for (int ks = 0; ks < kGqTile / 16; ++ks) {
#pragma unroll
      for (int dp = 0; dp < HS / 16; ++dp) {
        uint32_t vf[4];
        ldsm_x4_t(vf, vb + sw<HS>(ks * 16 + (((lane >> 3) & 1) << 3) + (lane & 7), dp * 2 + (lane >> 4)));
        pf_mma(o_acc[2 * dp], pfr[ks], vf[0], vf[1]);
        pf_mma(o_acc[2 * dp + 1], pfr[ks], vf[2], vf[3]);
      }
    }
    __syncwarp();   // the stage just read is refilled two iterations from now, by this warp's own cp.async
  }

  // ---- per-warp partial (m, l, acc) -> shared memory (each warp reuses ITS OWN stage memory), merge the 4 warps ----
  lA += __shfl_xor_sync(0xffffffffu, lA, 1); lA += __shfl_xor_sync(0xffffffffu, lA, 2);
  lB += __shfl_xor_sync(0xffffffffu, lB, 1); lB += __shfl_xor_sync(0xffffffffu, lB, 2);
  __syncwarp();
  float* const my_acc = reinterpret_cast<float*>(gq_smem + warp * kGqWarpBytes);   // [16][HS] fp32 = 8 KB
  if (g < nh) {
#pragma unroll
    for (int i = 0; i < HS / 8; ++i) *reinterpret_cast<float2*>(my_acc + g * HS + i * 8 + 2 * tq) = make_float2(o_acc[i][0], o_acc[i][1]);
    if (tq == 0) { s_m[warp][g] = mA; s_l[warp][g] = lA; }
  }
  if (g + 8 < nh) {
#pragma unroll
    for (int i = 0; i < HS / 8; ++i) *reinterpret_cast<float2*>(my_acc + (g + 8) * HS + i * 8 + 2 * tq) = make_float2(o_acc[i][2], o_acc[i][3]);
    if (tq == 0) { s_m[warp][g + 8] = mB; s_l[warp][g + 8] = lB; }
  }
  __syncthreads();

  const int H = p.n_head;
  auto warp_acc = [&](int w) -> const float* { return reinterpret_cast<const float*>(gq_smem + w * kGqWarpBytes); };
  for (int h = 0; h < nh; ++h) {
    const int head = gi * p.qpk + h;
    const float mn = fmaxf(fmaxf(s_m[0][h], s_m[1][h]), fmaxf(s_m[2][h], s_m[3][h]));
    float c[4], lt = 0.f;
#pragma unroll
    for (int w = 0; w < 4; ++w) { c[w] = mn == -INFINITY ? 0.f : exp2f(s_m[w][h] - mn); lt += s_l[w][h] * c[w]; }
    {
      const int d = tid;
      float a = 0.f;
#pragma unroll
      for (int w = 0; w < 4; ++w) a += c[w] == 0.f ? 0.f : warp_acc(w)[h * HS + d] * c[w];   // idle warps hold stale bytes
      if (p.nsplit == 1) p.out[((int64_t)b * H + head) * HS + d] = __float2bfloat16_rn(a / lt);
      else if (CLUSTER) reinterpret_cast<float*>(gq_smem)[h * HS + d] = a;   // element (h, d) of warp 0's region: only this thread read it
      else __stcg(p.ws_acc + (((int64_t)b * H + head) * p.nsplit + split) * HS + d, a);
    }
    if (p.nsplit > 1 && tid == 0) {
      if (CLUSTER) { s_pm[h] = mn; s_pl[h] = lt; }
      else {
        float* ml = p.ws_ml + (((int64_t)b * H + head) * p.nsplit + split) * 2;
        __stcg(ml, mn);
        __stcg(ml + 1, lt);
      }
    }
  }
  if (p.nsplit == 1) return;
  if constexpr (CLUSTER) {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (split == 0) {
      const int NS = p.nsplit;
      const uint32_t part_a = pf_smem(gq_smem), pm_a = pf_smem(&s_pm[0]), pl_a = pf_smem(&s_pl[0]);
      auto remote = [](uint32_t addr, int rank) -> float {
        uint32_t ra;
        float v;
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(addr), "r"(rank));
        asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(ra) : "memory");
        return v;
      };
      for (int i = tid; i < nh * HS; i += 128) {
        const int h = i / HS;
        float ms[8], ls[8], vs[8];
#pragma unroll
        for (int r = 0; r < 8; ++r) {
          ms[r] = r < NS ? remote(pm_a + h * 4, r) : -INFINITY;
          ls[r] = r < NS ? remote(pl_a + h * 4, r) : 0.f;
          vs[r] = r < NS ? remote(part_a + i * 4, r) : 0.f;
        }
        float mn = -INFINITY, lt = 0.f, a = 0.f;
#pragma unroll
        for (int r = 0; r < 8; ++r) mn = fmaxf(mn, ms[r]);
#pragma unroll
        for (int r = 0; r < 8; ++r) lt += ms[r] == -INFINITY ? 0.f : ls[r] * exp2f(ms[r] - mn);
        const float inv = 1.f / lt;
#pragma unroll
        for (int r = 0; r < 8; ++r) a = fmaf(vs[r], ms[r] == -INFINITY ? 0.f : exp2f(ms[r] - mn) * inv, a);
        p.out[((int64_t)b * H + gi * p.qpk + h) * HS + (i % HS)] = __float2bfloat16_rn(a);
      }
    }
    // no CTA leaves (and frees its shared memory) before rank 0 has read it
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    return;
  }

  // ---- last CTA of this (batch, group) merges the splits (protocol of attn_body.cuh) -----------------------------
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    unsigned int* cnt = p.ws_cnt + (int64_t)b * p.n_groups + gi;
    const unsigned int old = atomicAdd(cnt, 1u);
    s_last = old == (unsigned int)(p.nsplit - 1);
    if (s_last) *cnt = 0u;  // self-cleaning
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  float* s_w = reinterpret_cast<float*>(gq_smem);     // [nh][nsplit] weights, then the raw maxima
  float* s_mraw = s_w + kGqRows * 64;
  const int NS = p.nsplit;
  for (int i = tid; i < nh * NS; i += 128) {
    const int h = i / NS, sp = i % NS;
    const float* ml = p.ws_ml + (((int64_t)b * H + gi * p.qpk + h) * NS + sp) * 2;
    s_mraw[i] = __ldcg(ml);
    s_w[i] = __ldcg(ml + 1);
  }
  __syncthreads();
  if (tid < nh) {
    float mn = -INFINITY, lt = 0.f;
    for (int sp = 0; sp < NS; ++sp) mn = fmaxf(mn, s_mraw[tid * NS + sp]);
    for (int sp = 0; sp < NS; ++sp) lt += s_mraw[tid * NS + sp] == -INFINITY ? 0.f : s_w[tid * NS + sp] * exp2f(s_mraw[tid * NS + sp] - mn);
    const float inv = 1.f / lt;
    for (int sp = 0; sp < NS; ++sp) s_w[tid * NS + sp] = s_mraw[tid * NS + sp] == -INFINITY ? 0.f : exp2f(s_mraw[tid * NS + sp] - mn) * inv;
  }
  __syncthreads();
  for (int i = tid; i < nh * (HS / 4); i += 128) {   // one (head, 4 dims) item per thread, eight split loads in flight
    const int h = i / (HS / 4), d = (i % (HS / 4)) * 4;
    const int head = gi * p.qpk + h;
    const float* src = p.ws_acc + ((int64_t)b * H + head) * NS * HS + d;
    const float* w = s_w + h * NS;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int sp = 0; sp < NS; sp += 8) {
      float4 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = sp + u < NS ? __ldcg(reinterpret_cast<const float4*>(src + (int64_t)(sp + u) * HS)) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const float wu = sp + u < NS ? w[sp + u] : 0.f;
        a.x = fmaf(v[u].x, wu, a.x); a.y = fmaf(v[u].y, wu, a.y); a.z = fmaf(v[u].z, wu, a.z); a.w = fmaf(v[u].w, wu, a.w);
      }
    }
    *reinterpret_cast<uint2*>(p.out + ((int64_t)b * H + head) * HS + d) = make_uint2(pack_bf16(a.x, a.y), pack_bf16(a.z, a.w));
  }
}

}  // namespace

bool attn_decode_gqa_supported(int head_size, int qpk) { return head_size == kGqHS && qpk >= 2 && qpk <= kGqRows; }

cudaError_t launch_attn_decode_gqa(const AttnParams& p, dim3 grid, cudaStream_t st) {
  const bool cluster = grid.x >= 2 && grid.x <= 8;
  auto kernel = cluster ? attn_decode_gqa_kernel<true> : attn_decode_gqa_kernel<false>;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kGqSmem);
  if (e != cudaSuccess) return e;
  if (!cluster) return launch(kernel, grid, dim3(128), (size_t)kGqSmem, st, p);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(128);
  cfg.dynamicSmemBytes = (size_t)kGqSmem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = grid.x;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = g_pdl.load(std::memory_order_relaxed) ? 2 : 1;
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return cudaLaunchKernelEx(&cfg, kernel, p);
}

}  // namespace b200
