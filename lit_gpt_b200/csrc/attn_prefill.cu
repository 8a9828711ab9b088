// Prefill attention over the GQA-native KV cache (T > 1 query tokens), replacing the library SDPA call.
//
// Replaces CausalSelfAttention.scaled_dot_product_attention for the cached multi-token call
// (lit_gpt/model.py:646-665): q (B, H, T, hs) already roped, keys/values read from the caches
// (B, G, Tmax, hs) that b200_rope_kv_append has just extended.  Query row t sees cache positions j <= pos[t]
// -- exactly the rows `mask_cache.index_select(2, input_pos)` selects (model.py:509, 802-804) -- so chunked
// prefill (pos not starting at 0) works; positions past pos[t] are never read.  Softmax in fp32 with a running
// maximum (online softmax), probabilities rounded to bf16 for the P.V product like the fused SDPA back ends.
//
// One CTA = 128 query rows of one head (8 warps x 16 rows), key tiles of 64 positions double-buffered in shared
// memory by cp.async; S = Q K^T and O += P V on mma.sync.m16n8k16 (bf16 -> fp32) with ldmatrix operand loads from
// XOR-swizzled tiles.  The heads of one query group run in adjacent CTAs so the K/V tiles are shared through L2.
// Output is written as (B, T, H * hs), the layout attn.proj consumes (model.py:653).
#include "mma_tile.cuh"

namespace b200 {

namespace {

constexpr int kPfBM = 128;   // query rows per CTA
constexpr int kPfBN = 64;    // keys per tile
constexpr int kPfThreads = 256;

template <int HS, typename PosT>
__global__ void __launch_bounds__(kPfThreads, 1)
attn_prefill_kernel(const __nv_bfloat16* __restrict__ q, const __nv_bfloat16* __restrict__ kc, const __nv_bfloat16* __restrict__ vc,
                    const PosT* __restrict__ pos, __nv_bfloat16* __restrict__ y, int T, int n_head, int n_groups, int max_seq,
                    float scale_log2) {
  extern __shared__ __align__(128) uint8_t pf_smem_raw[];
  constexpr int CH = HS / 8;                         // 16-byte chunks per row
  constexpr int QB = kPfBM * HS * 2, KB = kPfBN * HS * 2;
  uint8_t* const sQ = pf_smem_raw;                   // [128][HS]
  uint8_t* const sK = sQ + QB;                       // [2][64][HS]
  uint8_t* const sV = sK + 2 * KB;                   // [2][64][HS]
  __shared__ int s_maxpos[kPfThreads / 32];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, tq = lane & 3;
  const int mt = (int)gridDim.x - 1 - (int)blockIdx.x;     // long (late) query tiles are scheduled first
  const int h = blockIdx.y, b = blockIdx.z;
  const int grp = h / (n_head / n_groups);
  const int r0 = mt * kPfBM;
  griddep_launch();
  griddep_wait();

  const __nv_bfloat16* qh = q + ((int64_t)b * n_head + h) * (int64_t)T * HS;
  const __nv_bfloat16* kg = kc + ((int64_t)b * n_groups + grp) * (int64_t)max_seq * HS;
  const __nv_bfloat16* vg = vc + ((int64_t)b * n_groups + grp) * (int64_t)max_seq * HS;

  // this thread's two query rows and their causal limits; the CTA's key range is [0, max limit]
  const int rowA = r0 + warp * 16 + g, rowB = rowA + 8;
  auto clampp = [&](int r) -> int {
    if (r >= T) return -1;
    int64_t v = (int64_t)pos[r];
    return (int)(v < 0 ? 0 : (v >= max_seq ? max_seq - 1 : v));
  };
  const int limA = clampp(rowA), limB = clampp(rowB);
  int wl = max(limA, limB);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) wl = max(wl, __shfl_xor_sync(0xffffffffu, wl, o));
  if (lane == 0) s_maxpos[warp] = wl;

  // Q tile -> shared (rows past T are zero-filled)
  for (int i = tid; i < kPfBM * CH; i += kPfThreads) {
    const int r = i / CH, c = i % CH;
    const bool ok = r0 + r < T;
    pf_cp16(pf_smem(sQ) + sw<HS>(r, c), qh + (int64_t)(ok ? r0 + r : 0) * HS + c * 8, ok);
  }
  pf_commit();
  __syncthreads();
  int cta_lim = -1;
#pragma unroll
  for (int w = 0; w < kPfThreads / 32; ++w) cta_lim = max(cta_lim, s_maxpos[w]);
  const int ntiles = (cta_lim + kPfBN) / kPfBN;      // cta_lim >= 0 because r0 < T

  auto load_kv = [&](int tile, int buf) {
    const int j0 = tile * kPfBN;
    for (int i = tid; i < kPfBN * CH; i += kPfThreads) {
      const int r = i / CH, c = i % CH;
      const bool ok = j0 + r <= cta_lim;
      const int64_t off = (int64_t)(ok ? j0 + r : 0) * HS + c * 8;
      pf_cp16(pf_smem(sK) + buf * KB + sw<HS>(r, c), kg + off, ok);
      pf_cp16(pf_smem(sV) + buf * KB + sw<HS>(r, c), vg + off, ok);
    }
    pf_commit();
  };
  load_kv(0, 0);

  float o_acc[HS / 8][4];
#pragma unroll
  for (int i = 0; i < HS / 8; ++i) o_acc[i][0] = o_acc[i][1] = o_acc[i][2] = o_acc[i][3] = 0.f;
  float mA = -INFINITY, mB = -INFINITY, lA = 0.f, lB = 0.f;

  // Q fragments stay in registers for the whole key loop
  pf_wait<1>();
  __syncthreads();
  uint32_t qf[HS / 16][4];
#pragma unroll
  for (int ks = 0; ks < HS / 16; ++ks)
    ldsm_x4(qf[ks], pf_smem(sQ) + sw<HS>(warp * 16 + (lane & 15), ks * 2 + (lane >> 4)));

  const int warp_lim = wl;
  for (int tile = 0; tile < ntiles; ++tile) {
    const int buf = tile & 1;
    if (tile + 1 < ntiles) { load_kv(tile + 1, buf ^ 1); pf_wait<1>(); } else { pf_wait<0>(); }
    __syncthreads();
    const int j0 = tile * kPfBN;
    if (j0 <= warp_lim) {   // warp-uniform: tiles entirely above this warp's rows contribute nothing
      float s[kPfBN / 8][4];
#pragma unroll
      for (int i = 0; i < kPfBN / 8; ++i) s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f;
      const uint32_t kb = pf_smem(sK) + buf * KB, vb = pf_smem(sV) + buf * KB;
#pragma unroll
      for (int ks = 0; ks < HS / 16; ++ks) {
#pragma unroll
        for (int np = 0; np < kPfBN / 16; ++np) {
          uint32_t kf[4];
          ldsm_x4(kf, kb + sw<HS>(np * 16 + ((lane >> 4) << 3) + (lane & 7), ks * 2 + ((lane >> 3) & 1)));
          pf_mma(s[2 * np], qf[ks], kf[0], kf[1]);
          pf_mma(s[2 * np + 1], qf[ks], kf[2], kf[3]);
        }
      }
      // causal limit + running maximum
      const bool need_mask = j0 + kPfBN - 1 > min(limA < 0 ? 0x7fffffff : limA, limB < 0 ? 0x7fffffff : limB) || limA < 0 || limB < 0;
      float tmA = -INFINITY, tmB = -INFINITY;
#pragma unroll
      for (int i = 0; i < kPfBN / 8; ++i) {
        if (need_mask) {
          const int j = j0 + i * 8 + 2 * tq;
          if (j > limA) s[i][0] = -INFINITY;
          if (j + 1 > limA) s[i][1] = -INFINITY;
          if (j > limB) s[i][2] = -INFINITY;
          if (j + 1 > limB) s[i][3] = -INFINITY;
        }
        tmA = fmaxf(tmA, fmaxf(s[i][0], s[i][1]));
        tmB = fmaxf(tmB, fmaxf(s[i][2], s[i][3]));
      }
      tmA = fmaxf(tmA, __shfl_xor_sync(0xffffffffu, tmA, 1)); tmA = fmaxf(tmA, __shfl_xor_sync(0xffffffffu, tmA, 2));
      tmB = fmaxf(tmB, __shfl_xor_sync(0xffffffffu, tmB, 1)); tmB = fmaxf(tmB, __shfl_xor_sync(0xffffffffu, tmB, 2));
      const float nmA = fmaxf(mA, tmA), nmB = fmaxf(mB, tmB);
      // rows with nothing visible yet keep m = -inf: use 0 as the exponent origin so that exp2(-inf - 0) = 0
      const float oA = nmA == -INFINITY ? 0.f : nmA * scale_log2, oB = nmB == -INFINITY ? 0.f : nmB * scale_log2;
      const float cA = mA == -INFINITY ? 0.f : exp2f(mA * scale_log2 - oA), cB = mB == -INFINITY ? 0.f : exp2f(mB * scale_log2 - oB);
      mA = nmA; mB = nmB;
      float rsA = 0.f, rsB = 0.f;
      uint32_t pfr[kPfBN / 16][4];
#pragma unroll
      for (int i = 0; i < kPfBN / 8; ++i) {
        const float p0 = exp2f(fmaf(s[i][0], scale_log2, -oA)), p1 = exp2f(fmaf(s[i][1], scale_log2, -oA));
        const float p2 = exp2f(fmaf(s[i][2], scale_log2, -oB)), p3 = exp2f(fmaf(s[i][3], scale_log2, -oB));
        rsA += p0 + p1; rsB += p2 + p3;
        pfr[i >> 1][(i & 1) * 2 + 0] = pack_bf16(p0, p1);
        pfr[i >> 1][(i & 1) * 2 + 1] = pack_bf16(p2, p3);
      }
      lA = lA * cA + rsA; lB = lB * cB + rsB;
#pragma unroll
      for (int i = 0; i < HS / 8; ++i) { o_acc[i][0] *= cA; o_acc[i][1] *= cA; o_acc[i][2] *= cB; o_acc[i][3] *= cB; }
#pragma unroll
      for (int ks = 0; ks < kPfBN / 16; ++ks) {
#pragma unroll
        for (int dp = 0; dp < HS / 16; ++dp) {
          uint32_t vf[4];
          ldsm_x4_t(vf, vb + sw<HS>(ks * 16 + (((lane >> 3) & 1) << 3) + (lane & 7), dp * 2 + (lane >> 4)));
          pf_mma(o_acc[2 * dp], pfr[ks], vf[0], vf[1]);
          pf_mma(o_acc[2 * dp + 1], pfr[ks], vf[2], vf[3]);
        }
      }
    }
    __syncthreads();   // the buffer just read is refilled by the next iteration's prefetch
  }

  lA += __shfl_xor_sync(0xffffffffu, lA, 1); lA += __shfl_xor_sync(0xffffffffu, lA, 2);
  lB += __shfl_xor_sync(0xffffffffu, lB, 1); lB += __shfl_xor_sync(0xffffffffu, lB, 2);
  const float iA = lA > 0.f ? 1.f / lA : 0.f, iB = lB > 0.f ? 1.f / lB : 0.f;
  const int64_t ystride = (int64_t)n_head * HS;
  if (rowA < T) {
    __nv_bfloat16* yr = y + ((int64_t)b * T + rowA) * ystride + (int64_t)h * HS;
#pragma unroll
    for (int i = 0; i < HS / 8; ++i) *reinterpret_cast<uint32_t*>(yr + i * 8 + 2 * tq) = pack_bf16(o_acc[i][0] * iA, o_acc[i][1] * iA);
  }
  if (rowB < T) {
    __nv_bfloat16* yr = y + ((int64_t)b * T + rowB) * ystride + (int64_t)h * HS;
#pragma unroll
    for (int i = 0; i < HS / 8; ++i) *reinterpret_cast<uint32_t*>(yr + i * 8 + 2 * tq) = pack_bf16(o_acc[i][2] * iB, o_acc[i][3] * iB);
  }
}

template <int HS, typename PosT>
cudaError_t launch_prefill(const void* q, const void* kc, const void* vc, const void* pos, void* y, int B, int T, int H, int G, int max_seq,
                           float scale, cudaStream_t st) {
  const size_t smem = (size_t)(kPfBM + 4 * kPfBN) * HS * 2;
  static PerDeviceOnce once;
  if (once.need()) {
    cudaError_t e = cudaFuncSetAttribute(attn_prefill_kernel<HS, PosT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    once.done();
  }
  dim3 grid((T + kPfBM - 1) / kPfBM, H, B);
  return launch(attn_prefill_kernel<HS, PosT>, grid, dim3(kPfThreads), smem, st, (const __nv_bfloat16*)q, (const __nv_bfloat16*)kc,
                (const __nv_bfloat16*)vc, (const PosT*)pos, (__nv_bfloat16*)y, T, H, G, max_seq, scale * 1.4426950408889634f);
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" int b200_attn_prefill(const void* q, const void* k_cache, const void* v_cache, const void* pos, int pos_is_int64, void* y,
                                 int batch, int T, int n_head, int n_groups, int head_size, int max_seq, float scale,
                                 b200_stream_t stream) {
  B200_CHECK_ARG(q && k_cache && v_cache && pos && y, "b200_attn_prefill: null pointer");
  B200_CHECK_ARG(batch > 0 && T > 0 && n_head > 0 && n_groups > 0 && max_seq > 0, "b200_attn_prefill: bad shape");
  B200_CHECK_ARG(n_head % n_groups == 0, "b200_attn_prefill: n_head not divisible by n_groups");
  B200_CHECK_ARG(head_size == 64 || head_size == 128, "b200_attn_prefill: head_size must be 64 or 128 (got %d)", head_size);
  B200_CHECK_ARG(n_head <= 65535 && batch <= 65535, "b200_attn_prefill: grid too large");
  cudaStream_t st = (cudaStream_t)stream;
  if (head_size == 128) {
    if (pos_is_int64) B200_CUDA((launch_prefill<128, int64_t>(q, k_cache, v_cache, pos, y, batch, T, n_head, n_groups, max_seq, scale, st)));
    else B200_CUDA((launch_prefill<128, int32_t>(q, k_cache, v_cache, pos, y, batch, T, n_head, n_groups, max_seq, scale, st)));
  } else {
    if (pos_is_int64) B200_CUDA((launch_prefill<64, int64_t>(q, k_cache, v_cache, pos, y, batch, T, n_head, n_groups, max_seq, scale, st)));
    else B200_CUDA((launch_prefill<64, int32_t>(q, k_cache, v_cache, pos, y, batch, T, n_head, n_groups, max_seq, scale, st)));
  }
  return B200_OK;
}
