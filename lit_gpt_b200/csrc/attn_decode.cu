// Fused RoPE + KV-cache append + split-K flash-decode attention for T = 1.
//
// Replaces, for one decode step, lit_gpt/model.py:624-653: the qkv view/permute/split (624-628), K/V expand
// for GQA (633-635), apply_rope on q and k (641-644 -> 767-773), KVCache.forward index_copy_ (649 -> 788-795)
// and F.scaled_dot_product_attention over the FULL max_seq cache with a boolean mask row (651 -> 658-665).
// Here the cache is GQA-native (n_query_groups heads), only the filled positions 0..pos are read,
// and the q_per_kv query heads of a group share each K/V row from registers.
//
// HBM-bound on the KV stream (bytes = 2 * groups * ctx * head_size * 2 B per layer): rows are read with
// coalesced 8/16-byte loads, LPK lanes per key, online softmax in fp32 per lane group; lane groups, warps and
// splits are merged with the usual (m, l, acc) rescaling, the last CTA of a (batch, group) to finish doing the
// final merge (deterministic order), so the whole step is ONE launch.
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
};

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
template <int HS, int QPK, int DPL>
__global__ void __launch_bounds__(128) attn_decode_kernel(const AttnParams p) {
  constexpr int LPK = HS / DPL, GPW = 32 / LPK, SLOTS = 4 * GPW, NR = DPL / 2;
  __shared__ float s_q[QPK][HS];
  __shared__ float s_k[HS], s_v[HS], s_tmp[HS];
  __shared__ float s_acc[4][QPK][HS];
  __shared__ float s_m[4][QPK], s_l[4][QPK];
  __shared__ unsigned int s_last;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int split = blockIdx.x, b = blockIdx.z;
  const int gi = blockIdx.y / p.head_chunks, hc = blockIdx.y % p.head_chunks;
  const int h0 = hc * QPK;                                     // first query head (within the group) of this CTA
  const int nh = min(QPK, p.qpk - h0);                         // live heads
  griddep_launch();
  griddep_wait();

  int64_t pos = p.pos_i64 ? *reinterpret_cast<const int64_t*>(p.pos) : (int64_t)*reinterpret_cast<const int32_t*>(p.pos);
  pos = pos < 0 ? 0 : (pos >= p.max_seq ? p.max_seq - 1 : pos);  // host validates; never index out of bounds
  const bool last_split = split == p.nsplit - 1;

  // ---- q (and, in the last split, the new k / v): load, RoPE, keep as bf16-exact floats in smem -----------
  const __nv_bfloat16* grp = p.qkv + ((int64_t)b * p.n_groups + gi) * (p.qpk + 2) * HS;
  for (int h = 0; h < nh; ++h) {
    for (int d = tid; d < HS; d += 128) s_tmp[d] = __bfloat162float(grp[(h0 + h) * HS + d]);
    __syncthreads();
    for (int d = tid; d < HS; d += 128) s_q[h][d] = rope_elem(s_tmp, d, p.rope_n, p.cos, p.sin, p.table_bf16, p.table_presel ? 0 : pos);
    __syncthreads();
  }
  if (last_split) {
    for (int d = tid; d < HS; d += 128) { s_tmp[d] = __bfloat162float(grp[p.qpk * HS + d]); s_v[d] = __bfloat162float(grp[(p.qpk + 1) * HS + d]); }
    __syncthreads();
    for (int d = tid; d < HS; d += 128) s_k[d] = rope_elem(s_tmp, d, p.rope_n, p.cos, p.sin, p.table_bf16, p.table_presel ? 0 : pos);
    __syncthreads();
    if (hc == 0) {  // exactly one CTA per (batch, group) appends to the cache (KVCache.forward, model.py:793-794)
      const int64_t off = (((int64_t)b * p.n_groups + gi) * p.max_seq + pos) * HS;
      for (int d = tid; d < HS; d += 128) { p.kc[off + d] = __float2bfloat16_rn(s_k[d]); p.vc[off + d] = __float2bfloat16_rn(s_v[d]); }
    }
  }

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
  for (int64_t base = kbeg + warp * GPW; base < kend; base += 2 * SLOTS) {  // warp-uniform trip count
    const int64_t ka = base + gidx, kb = ka + SLOTS;
    const bool va = ka < kend, vb = kb < kend;
    uint32_t kra[NR], vra[NR], krb[NR], vrb[NR];
#pragma unroll
    for (int i = 0; i < NR; ++i) kra[i] = vra[i] = krb[i] = vrb[i] = 0u;
    if (va) { load_row<NR>(kra, kbase + ka * HS); load_row<NR>(vra, vbase + ka * HS); }
    if (vb) { load_row<NR>(krb, kbase + kb * HS); load_row<NR>(vrb, vbase + kb * HS); }
    update(kra, vra, va);
    if (base + SLOTS < kend) update(krb, vrb, vb);
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
  __syncthreads();

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
      else __stcg(p.ws_acc + (((int64_t)b * H + head) * p.nsplit + split) * HS + d, a);
    }
    if (p.nsplit > 1 && tid == 0) {
      float* ml = p.ws_ml + (((int64_t)b * H + head) * p.nsplit + split) * 2;
      __stcg(ml, mn);
      __stcg(ml + 1, lt);
    }
  }
  if (p.nsplit == 1) return;

  // ---- last CTA of this (batch, group, head chunk) merges the splits ---------------------------------------
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    unsigned int* cnt = p.ws_cnt + (int64_t)b * p.n_groups * p.head_chunks + blockIdx.y;
    const unsigned int old = atomicAdd(cnt, 1u);
    s_last = old == (unsigned int)(p.nsplit - 1);
    if (s_last) *cnt = 0u;  // self-cleaning
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  for (int h = 0; h < nh; ++h) {
    const int head = gi * p.qpk + h0 + h;
    const float* ml = p.ws_ml + ((int64_t)b * H + head) * p.nsplit * 2;
    float mn = -INFINITY;
    for (int s = 0; s < p.nsplit; ++s) mn = fmaxf(mn, __ldcg(ml + 2 * s));
    float lt = 0.f;
    for (int s = 0; s < p.nsplit; ++s) lt += __ldcg(ml + 2 * s + 1) * exp2f(__ldcg(ml + 2 * s) - mn);
    for (int d = tid; d < HS; d += 128) {
      float a = 0.f;
      for (int s = 0; s < p.nsplit; ++s)
        a += __ldcg(p.ws_acc + (((int64_t)b * H + head) * p.nsplit + s) * HS + d) * exp2f(__ldcg(ml + 2 * s) - mn);
      p.out[((int64_t)b * H + head) * HS + d] = __float2bfloat16_rn(a / lt);
    }
  }
}

// ---- prefill helper: RoPE + cache write for T > 1 (lit_gpt/model.py:624-649) --------------------------------
// grid (T, n_groups, batch); q_out is (batch, n_head, T, hs) so torch SDPA / the prefill attention can consume it.
template <typename PosT>
__global__ void __launch_bounds__(128) rope_kv_append_kernel(const __nv_bfloat16* __restrict__ qkv, const void* cos, const void* sin,
                                                             int table_bf16, int table_presel, int rope_n, const PosT* __restrict__ pos,
                                                             __nv_bfloat16* __restrict__ kc, __nv_bfloat16* __restrict__ vc,
                                                             __nv_bfloat16* __restrict__ q_out, int T, int n_head, int n_groups,
                                                             int hs, int max_seq) {
  extern __shared__ float s_x[];  // one head vector
  griddep_launch();
  griddep_wait();
  const int t = blockIdx.x, gi = blockIdx.y, b = blockIdx.z, tid = threadIdx.x;
  const int qpk = n_head / n_groups;
  int64_t ps = (int64_t)pos[t];
  ps = ps < 0 ? 0 : (ps >= max_seq ? max_seq - 1 : ps);
  const __nv_bfloat16* grp = qkv + (((int64_t)b * T + t) * n_groups + gi) * (qpk + 2) * hs;
  for (int h = 0; h <= qpk; ++h) {  // q heads then k
    for (int d = tid; d < hs; d += 128) s_x[d] = __bfloat162float(grp[h * hs + d]);
    __syncthreads();
    for (int d = tid; d < hs; d += 128) {
      const float r = rope_elem(s_x, d, rope_n, cos, sin, table_bf16, table_presel ? (int64_t)t : ps);
      if (h < qpk) q_out[(((int64_t)b * n_head + gi * qpk + h) * T + t) * hs + d] = __float2bfloat16_rn(r);
      else kc[(((int64_t)b * n_groups + gi) * max_seq + ps) * hs + d] = __float2bfloat16_rn(r);
    }
    __syncthreads();
  }
  for (int d = tid; d < hs; d += 128) vc[(((int64_t)b * n_groups + gi) * max_seq + ps) * hs + d] = grp[(qpk + 1) * hs + d];
}

template <int HS>
static cudaError_t launch_attn(const AttnParams& p, int qpk_t, dim3 grid, cudaStream_t st) {
  switch (qpk_t) {
    case 1: return launch(attn_decode_kernel<HS, 1, 16>, grid, dim3(128), 0, st, p);
    case 2: return launch(attn_decode_kernel<HS, 2, 16>, grid, dim3(128), 0, st, p);
    case 4: return launch(attn_decode_kernel<HS, 4, 8>, grid, dim3(128), 0, st, p);
    default: return launch(attn_decode_kernel<HS, 8, 4>, grid, dim3(128), 0, st, p);
  }
}

}  // namespace b200

using namespace b200;

extern "C" {

size_t b200_attn_decode_workspace_bytes(int batch, int n_head, int head_size, int max_splits) {
  size_t acc = (size_t)batch * n_head * max_splits * head_size * sizeof(float);
  size_t ml = (size_t)batch * n_head * max_splits * 2 * sizeof(float);
  size_t cnt = (size_t)batch * n_head * sizeof(unsigned int);
  return acc + ml + cnt + 256;
}

int b200_attn_decode(const b200_attn_decode_args* a, b200_stream_t stream) {
  B200_CHECK_ARG(a != nullptr, "b200_attn_decode: null args");
  B200_CHECK_ARG(a->qkv && a->cos && a->sin && a->pos && a->k_cache && a->v_cache && a->out, "b200_attn_decode: null pointer");
  B200_CHECK_ARG(a->batch > 0 && a->n_head > 0 && a->n_groups > 0 && a->max_seq > 0, "b200_attn_decode: bad shape");
  B200_CHECK_ARG(a->n_head % a->n_groups == 0, "b200_attn_decode: n_head %d not divisible by n_groups %d", a->n_head, a->n_groups);
  if (a->head_size != 64 && a->head_size != 128) {
    set_error("b200_attn_decode: head_size %d not supported by the fused kernel (64 or 128)", a->head_size);
    return B200_ERR_UNSUPPORTED;
  }
  B200_CHECK_ARG(a->rope_n_elem >= 0 && a->rope_n_elem <= a->head_size && a->rope_n_elem % 2 == 0, "b200_attn_decode: bad rope_n_elem %d", a->rope_n_elem);
  B200_CHECK_ARG(a->table_dtype == B200_F32 || a->table_dtype == B200_BF16, "b200_attn_decode: cos/sin must be fp32 or bf16");
  const int qpk = a->n_head / a->n_groups;
  const int qpk_t = qpk == 1 ? 1 : (qpk == 2 ? 2 : (qpk <= 4 ? 4 : 8));
  const int head_chunks = (qpk + qpk_t - 1) / qpk_t;
  int nsplit = a->num_splits;
  if (nsplit <= 0) {
    const int ctas = a->batch * a->n_groups * head_chunks;
    nsplit = (2 * sm_count() + ctas - 1) / ctas;
    const int by_len = (a->max_seq + 63) / 64;
    if (nsplit > by_len) nsplit = by_len;
  }
  if (nsplit > 64) nsplit = 64;
  if (nsplit < 1) nsplit = 1;
  size_t need = b200_attn_decode_workspace_bytes(a->batch, a->n_head, a->head_size, nsplit);
  B200_CHECK_ARG(nsplit == 1 || (a->workspace && a->workspace_bytes >= need), "b200_attn_decode: workspace too small (%zu < %zu)", a->workspace_bytes, need);

  AttnParams p{};
  p.qkv = (const __nv_bfloat16*)a->qkv;
  p.cos = a->cos; p.sin = a->sin; p.table_bf16 = a->table_dtype == B200_BF16; p.table_presel = a->table_preselected; p.rope_n = a->rope_n_elem;
  p.pos = a->pos; p.pos_i64 = a->pos_is_int64;
  p.kc = (__nv_bfloat16*)a->k_cache; p.vc = (__nv_bfloat16*)a->v_cache; p.out = (__nv_bfloat16*)a->out;
  p.batch = a->batch; p.n_head = a->n_head; p.n_groups = a->n_groups; p.hs = a->head_size; p.max_seq = a->max_seq; p.qpk = qpk;
  p.nsplit = nsplit; p.head_chunks = head_chunks;
  if (nsplit > 1) {
    uint8_t* ws = (uint8_t*)a->workspace;
    // counters first (must stay zero between launches; the kernel cleans up after itself)
    p.ws_cnt = (unsigned int*)ws;
    size_t off = (((size_t)a->batch * a->n_head * sizeof(unsigned int)) + 255) & ~(size_t)255;
    p.ws_ml = (float*)(ws + off);
    off += (size_t)a->batch * a->n_head * nsplit * 2 * sizeof(float);
    p.ws_acc = (float*)(ws + off);
  }
  p.scale_log2 = (1.0f / sqrtf((float)a->head_size)) * 1.4426950408889634f;
  dim3 grid(nsplit, a->n_groups * head_chunks, a->batch);
  cudaError_t e = a->head_size == 128 ? launch_attn<128>(p, qpk_t, grid, (cudaStream_t)stream) : launch_attn<64>(p, qpk_t, grid, (cudaStream_t)stream);
  if (e != cudaSuccess) {
    set_error("b200_attn_decode: launch failed: %s", cudaGetErrorString(e));
    return B200_ERR_CUDA;
  }
  return B200_OK;
}

int b200_rope_kv_append(const void* qkv, const void* cos, const void* sin, int table_dtype, int table_preselected, int rope_n_elem, const void* pos,
                        int pos_is_int64, void* k_cache, void* v_cache, void* q_out, int batch, int T, int n_head, int n_groups,
                        int head_size, int max_seq, b200_stream_t stream) {
  B200_CHECK_ARG(qkv && cos && sin && pos && k_cache && v_cache && q_out, "b200_rope_kv_append: null pointer");
  B200_CHECK_ARG(batch > 0 && T > 0 && n_head > 0 && n_groups > 0 && head_size > 0 && max_seq > 0, "b200_rope_kv_append: bad shape");
  B200_CHECK_ARG(n_head % n_groups == 0, "b200_rope_kv_append: n_head not divisible by n_groups");
  B200_CHECK_ARG(rope_n_elem >= 0 && rope_n_elem <= head_size && rope_n_elem % 2 == 0, "b200_rope_kv_append: bad rope_n_elem");
  B200_CHECK_ARG(table_dtype == B200_F32 || table_dtype == B200_BF16, "b200_rope_kv_append: cos/sin must be fp32 or bf16");
  B200_CHECK_ARG(T <= 65535 * 32 && n_groups <= 65535 && batch <= 65535, "b200_rope_kv_append: grid too large");
  dim3 grid(T, n_groups, batch);
  size_t smem = (size_t)head_size * sizeof(float);
  cudaStream_t st = (cudaStream_t)stream;
  const int tb = table_dtype == B200_BF16;
  if (pos_is_int64)
    B200_CUDA(launch(rope_kv_append_kernel<int64_t>, grid, dim3(128), smem, st, (const __nv_bfloat16*)qkv, cos, sin, tb, table_preselected, rope_n_elem, (const int64_t*)pos, (__nv_bfloat16*)k_cache, (__nv_bfloat16*)v_cache, (__nv_bfloat16*)q_out, T, n_head, n_groups, head_size, max_seq));
  else
    B200_CUDA(launch(rope_kv_append_kernel<int32_t>, grid, dim3(128), smem, st, (const __nv_bfloat16*)qkv, cos, sin, tb, table_preselected, rope_n_elem, (const int32_t*)pos, (__nv_bfloat16*)k_cache, (__nv_bfloat16*)v_cache, (__nv_bfloat16*)q_out, T, n_head, n_groups, head_size, max_seq));
  return B200_OK;
}

}  // extern "C"
