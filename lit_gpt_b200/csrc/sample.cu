// Fused on-device sampling for the decode loop: top-k truncation + temperature softmax + one multinomial draw, plus the
// loop bookkeeping (position increment, history append), as ONE launch that stays inside the decode CUDA graph.
//
// Replaces sample() / multinomial_num_samples_1() of generate/base.py:22-41 (topk + scatter into -inf, softmax of
// logits / temperature, and the draw argmax(probs / Exp(1)) that the reference itself uses under torch.compile,
// base.py:23-26) and the per-token bookkeeping of base.py:90-93.  Like the reference it reads batch row 0 only.
//   kept set   : the top_k largest logits; ties at the threshold are kept in index order until top_k are kept
//   p_i        : exp((l_i - max) / T) / sum over the kept set, fp32
//   token      : argmax_i p_i / e_i,  e_i ~ Exp(1) from Philox4x32-10(seed, offset, i) or from a caller-given noise row
//   T <= 0     : greedy (argmax of the logits, lowest index on ties)
// One CTA of 1024 threads; thread t owns the contiguous index range [t * chunk, (t + 1) * chunk) so that tie ranks follow
// index order; the threshold is found by an 8-bit radix select over order-preserving keys (4 histogram passes).
#include "common.cuh"

namespace b200 {

namespace {

__device__ __forceinline__ uint32_t order_key(float v) {   // ascending float order == ascending unsigned order
  const uint32_t u = __float_as_uint(v);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t (&out)[4]) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    c0 = hi1 ^ c1 ^ k0; c1 = lo1; c2 = hi0 ^ c3 ^ k1; c3 = lo0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

template <typename LT>
__device__ __forceinline__ float ld_logit(const LT* p, int64_t i) { return load_as_float<LT>(p, i); }

template <typename LT, typename TokT, typename PosT>
__global__ void __launch_bounds__(1024) sample_kernel(const LT* __restrict__ logits, int vocab, float temperature, int top_k,
                                                      int64_t* __restrict__ rng_state, const float* __restrict__ noise,
                                                      TokT* __restrict__ token_out, PosT* __restrict__ pos_inout,
                                                      TokT* __restrict__ history, int64_t history_len) {
  __shared__ uint32_t s_hist[256];
  __shared__ uint32_t s_bin, s_rem;
  __shared__ float s_f[32];
  __shared__ int s_i[32];
  __shared__ uint32_t s_scan[32];
  griddep_launch();
  griddep_wait();
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int chunk = (vocab + 1023) / 1024;
  const int i0 = min(vocab, tid * chunk), i1 = min(vocab, i0 + chunk);

  auto block_max = [&](float v) -> float {
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) s_f[warp] = v;
    __syncthreads();
    float r = s_f[lane];
    return warp_max(r);
  };
  auto block_sum = [&](float v) -> float {
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) s_f[warp] = v;
    __syncthreads();
    float r = s_f[lane];
    return warp_sum(r);
  };

  // ---- top-k threshold: radix select of the top_k-th largest key ----------------------------------------------
  const bool trunc = top_k > 0 && top_k < vocab && temperature > 0.f;
  uint32_t tkey = 0u;
  int keep_ties = 0, tie_base = 0;
  if (trunc) {
    uint32_t prefix = 0u, mask = 0u;
    int remaining = top_k;
    for (int pass = 0; pass < 4; ++pass) {
      const int shift = 24 - 8 * pass;
      if (tid < 256) s_hist[tid] = 0u;
      __syncthreads();
      for (int i = i0; i < i1; ++i) {
        const uint32_t key = order_key(ld_logit(logits, i));
        if ((key & mask) == prefix) atomicAdd(&s_hist[(key >> shift) & 255u], 1u);
      }
      __syncthreads();
      if (warp == 0) {  // bins from the top: the bin where the cumulative count reaches `remaining`
        uint32_t loc[8], tot = 0u;
#pragma unroll
        for (int j = 0; j < 8; ++j) { loc[j] = s_hist[255 - (lane * 8 + j)]; tot += loc[j]; }
        uint32_t inc = tot;  // inclusive scan over lanes (lane 0 = highest bins)
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
        const uint32_t exc = inc - tot;
        if (exc < (uint32_t)remaining && inc >= (uint32_t)remaining) {  // exactly one lane
          uint32_t cum = exc;
          int j = 0;
          for (; j < 8; ++j) { if (cum + loc[j] >= (uint32_t)remaining) break; cum += loc[j]; }
          s_bin = 255u - (uint32_t)(lane * 8 + j);
          s_rem = (uint32_t)remaining - cum;
        }
      }
      __syncthreads();
      prefix |= s_bin << shift;
      mask |= 0xFFu << shift;
      remaining = (int)s_rem;
      __syncthreads();
    }
    tkey = prefix;
    keep_ties = remaining;  // >= 1 elements equal to the threshold are kept, in index order
    int mine = 0;
    for (int i = i0; i < i1; ++i) mine += order_key(ld_logit(logits, i)) == tkey;
    // exclusive block scan of the per-thread tie counts (threads are in index order)
    int inc = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
    if (lane == 31) s_scan[warp] = (uint32_t)inc;
    __syncthreads();
    if (warp == 0) {
      int w = (int)s_scan[lane], winc = w;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, winc, o); if (lane >= o) winc += t; }
      s_scan[lane] = (uint32_t)(winc - w);
    }
    __syncthreads();
    tie_base = (int)s_scan[warp] + inc - mine;
  }

  // ---- max, normaliser, draw ---------------------------------------------------------------------------------------
  float lmax = -INFINITY;
  for (int i = i0; i < i1; ++i) lmax = fmaxf(lmax, ld_logit(logits, i));
  lmax = block_max(lmax);
  const bool greedy = !(temperature > 0.f);
  const float invT = greedy ? 0.f : 1.4426950408889634f / temperature;
  auto kept = [&](uint32_t key, int& tie_rank) -> bool {
    if (!trunc || key > tkey) return true;
    if (key < tkey) return false;
    return tie_rank++ < keep_ties;
  };
  float psum = 0.f;
  if (!greedy) {
    int tr = tie_base;
    for (int i = i0; i < i1; ++i) {
      const float l = ld_logit(logits, i);
      if (kept(order_key(l), tr)) psum += exp2f((l - lmax) * invT);
    }
    psum = block_sum(psum);
  }
  const uint64_t seed = rng_state ? (uint64_t)rng_state[0] : 0ull, offset = rng_state ? (uint64_t)rng_state[1] : 0ull;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  {
    int tr = tie_base;
    uint32_t rnd[4] = {0u, 0u, 0u, 0u};
    int have = -1;
    for (int i = i0; i < i1; ++i) {
      const float l = ld_logit(logits, i);
      float score;
      if (greedy) {
        score = l;
      } else {
        if (!kept(order_key(l), tr)) continue;
        float e;
        if (noise) {
          e = noise[i];
        } else {
          if (have != (i >> 2)) { have = i >> 2; philox4x32_10((uint32_t)have, 0u, (uint32_t)offset, (uint32_t)(offset >> 32), (uint32_t)seed, (uint32_t)(seed >> 32), rnd); }
          const float u = ((float)(rnd[i & 3] >> 8) + 0.5f) * (1.0f / 16777216.0f);   // (0, 1)
          e = -__logf(u);
        }
        score = exp2f((l - lmax) * invT) / psum / e;
      }
      if (score > best || bi == 0x7fffffff) { best = score; bi = i; }   // index order within the thread: first maximum wins
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (oi != 0x7fffffff && (bi == 0x7fffffff || ov > best || (ov == best && oi < bi))) { best = ov; bi = oi; }
  }
  __syncthreads();
  if (lane == 0) { s_f[warp] = best; s_i[warp] = bi; }
  __syncthreads();
  if (warp == 0) {
    best = s_f[lane];
    bi = s_i[lane];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (oi != 0x7fffffff && (bi == 0x7fffffff || ov > best || (ov == best && oi < bi))) { best = ov; bi = oi; }
    }
    if (lane == 0) {
      token_out[0] = (TokT)bi;
      if (rng_state) rng_state[1] = (int64_t)(offset + 1ull);
      if (pos_inout) {
        const int64_t np = (int64_t)pos_inout[0] + 1;
        pos_inout[0] = (PosT)np;
        if (history && np >= 0 && np < history_len) history[np] = (TokT)bi;
      }
    }
  }
}

template <typename LT>
cudaError_t launch_sample(const void* logits, int vocab, float temperature, int top_k, void* rng, const float* noise, void* tok, int tok64,
                          void* pos, int pos64, void* hist, int64_t hist_len, cudaStream_t st) {
  const LT* l = (const LT*)logits;
  dim3 grid(1), block(1024);
  if (tok64 && pos64) return launch(sample_kernel<LT, int64_t, int64_t>, grid, block, 0, st, l, vocab, temperature, top_k, (int64_t*)rng, noise, (int64_t*)tok, (int64_t*)pos, (int64_t*)hist, hist_len);
  if (tok64) return launch(sample_kernel<LT, int64_t, int32_t>, grid, block, 0, st, l, vocab, temperature, top_k, (int64_t*)rng, noise, (int64_t*)tok, (int32_t*)pos, (int64_t*)hist, hist_len);
  if (pos64) return launch(sample_kernel<LT, int32_t, int64_t>, grid, block, 0, st, l, vocab, temperature, top_k, (int64_t*)rng, noise, (int32_t*)tok, (int64_t*)pos, (int32_t*)hist, hist_len);
  return launch(sample_kernel<LT, int32_t, int32_t>, grid, block, 0, st, l, vocab, temperature, top_k, (int64_t*)rng, noise, (int32_t*)tok, (int32_t*)pos, (int32_t*)hist, hist_len);
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" int b200_sample(const void* logits, int logits_dtype, int64_t vocab, float temperature, int top_k, void* rng_state,
                           const void* noise, void* token_out, int token_is_int64, void* pos_inout, int pos_is_int64, void* history,
                           int64_t history_len, b200_stream_t stream) {
  B200_CHECK_ARG(logits && token_out, "b200_sample: null pointer");
  B200_CHECK_ARG(vocab > 0 && vocab < (1 << 30), "b200_sample: bad vocabulary size %lld", (long long)vocab);
  B200_CHECK_ARG(logits_dtype == B200_BF16 || logits_dtype == B200_F32, "b200_sample: logits must be bf16 or fp32");
  B200_CHECK_ARG(!(temperature > 0.f) || rng_state || noise, "b200_sample: temperature > 0 needs rng_state or noise");
  cudaStream_t st = (cudaStream_t)stream;
  if (logits_dtype == B200_BF16)
    B200_CUDA((launch_sample<__nv_bfloat16>(logits, (int)vocab, temperature, top_k, rng_state, (const float*)noise, token_out, token_is_int64, pos_inout, pos_is_int64, history, history_len, st)));
  else
    B200_CUDA((launch_sample<float>(logits, (int)vocab, temperature, top_k, rng_state, (const float*)noise, token_out, token_is_int64, pos_inout, pos_is_int64, history, history_len, st)));
  return B200_OK;
}
