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
#include "attn_body.cuh"

namespace b200 {

// HS head size, QPK query heads handled per CTA (shared K/V rows), DPL dims per lane (HS/DPL lanes per key)
template <int HS, int QPK, int DPL>
__global__ void __launch_bounds__(128) attn_decode_kernel(const AttnParams p) {
  __shared__ AttnSmem<HS, QPK> sm;
  griddep_launch();
  griddep_wait();
  attn_decode_item<HS, QPK, DPL>(p, sm, (int)threadIdx.x, (int)blockIdx.x, (int)blockIdx.y, (int)blockIdx.z, [] { __syncthreads(); });
}

struct CtaSync {
  __device__ __forceinline__ void operator()() const { __syncthreads(); }
};

// same body; the nsplit (<= 8) CTAs along x are one thread-block cluster and merge through distributed shared memory
template <int HS, int QPK, int DPL>
__global__ void __launch_bounds__(128) attn_decode_cluster_kernel(const AttnParams p) {
  __shared__ AttnSmem<HS, QPK> sm;
  griddep_launch();
  griddep_wait();
  attn_decode_item<HS, QPK, DPL, CtaSync, true>(p, sm, (int)threadIdx.x, (int)blockIdx.x, (int)blockIdx.y, (int)blockIdx.z, CtaSync{});
}

template <typename... KArgs, typename... Args>
static cudaError_t launch_cluster(void (*kernel)(KArgs...), dim3 grid, dim3 block, unsigned cluster_x, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = 0;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = cluster_x;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = g_pdl.load(std::memory_order_relaxed) ? 2 : 1;
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

template <int HS>
static cudaError_t launch_attn_cluster(const AttnParams& p, int qpk_t, dim3 grid, cudaStream_t st) {
  switch (qpk_t) {
    case 1: return launch_cluster(attn_decode_cluster_kernel<HS, 1, 16>, grid, dim3(128), grid.x, st, p);
    case 2: return launch_cluster(attn_decode_cluster_kernel<HS, 2, 16>, grid, dim3(128), grid.x, st, p);
    case 4: return launch_cluster(attn_decode_cluster_kernel<HS, 4, 8>, grid, dim3(128), grid.x, st, p);
    default: return launch_cluster(attn_decode_cluster_kernel<HS, 8, 4>, grid, dim3(128), grid.x, st, p);
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

static std::atomic<int> g_attn_force_scalar{0};

extern "C" {

int b200_attn_decode_tune(int force_scalar) { return g_attn_force_scalar.exchange(force_scalar ? 1 : 0); }

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
  // grouped / multi-query attention at head size 128: the heads of a group are the M rows of a tensor-core tile
  const bool force_scalar = g_attn_force_scalar.load(std::memory_order_relaxed) != 0;
  const bool gqa = attn_decode_gqa_supported(a->head_size, qpk) && !force_scalar;
  const int head_chunks = gqa ? 1 : (qpk + qpk_t - 1) / qpk_t;
  int nsplit = a->num_splits;
  if (nsplit <= 0) {
    const int ctas = a->batch * a->n_groups * head_chunks;
    // measured (tools/attn_bench.py --splits): the kernels are bound by dependent round trips, so short key ranges per
    // CTA win until the merge grows; about 3 CTAs per SM for the scalar kernel, 1 for the tensor-core one
    nsplit = ((gqa ? 1 : 3) * sm_count() + ctas - 1) / ctas;
    const int by_len = gqa ? (a->max_seq + 255) / 256 : (a->max_seq + 31) / 32;   // gqa: two 32-key tiles per warp at least
    if (nsplit > by_len) nsplit = by_len;
    // many sequences x heads over a long cache (13B, batch 16, 2048 positions: 640 CTAs of 1 MB each): three key ranges per (sequence,
    // head) still pay -- 136 -> 120 us per layer, 4.9 -> 5.6 TB/s (tools/attn_bench.py --splits) -- the last wave is shorter
    if (!gqa && nsplit < 3 && a->max_seq >= 3 * 512) nsplit = 3;
    // a cache of at most 128 positions: one CTA per (sequence, head) without the cluster merge -- 4.3 / 4.7 / 5.8 / 5.9 us at 27 / 60 / 100 /
    // 112 filled positions against 4.3 (5.7 clustered) / 5.5 / 5.9 / 5.9; the split wins from ~150 on (7.3 vs 6.2 us at 176):
    // tools/attn_bench.py --splits
    if (!gqa && a->max_seq <= 128) nsplit = 1;
    // scalar kernel: cap at the portable cluster size so that the splits merge through distributed shared memory (the
    // tensor-core kernel measured the same with 9..16 workspace-merged splits as with 8 clustered ones: not capped)
    if (!gqa && !force_scalar && nsplit > 8) nsplit = 8;
  }
  const bool cluster = !gqa && !force_scalar && nsplit >= 2 && nsplit <= 8;
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
  cudaError_t e = gqa ? launch_attn_decode_gqa(p, grid, (cudaStream_t)stream)
                  : cluster ? (a->head_size == 128 ? launch_attn_cluster<128>(p, qpk_t, grid, (cudaStream_t)stream) : launch_attn_cluster<64>(p, qpk_t, grid, (cudaStream_t)stream))
                      : (a->head_size == 128 ? launch_attn<128>(p, qpk_t, grid, (cudaStream_t)stream) : launch_attn<64>(p, qpk_t, grid, (cudaStream_t)stream));
  if (e != cudaSuccess) {
    set_error("b200_attn_decode: launch failed: %s", cudaGetErrorString(e));
    return B200_ERR_CUDA;
  }
  return B200_OK;
}

}  // extern "C"

namespace b200 {
int attn_fused_params(const b200_attn_decode_args* a, int nsplit, AttnParams* out) {
  if (!a || !a->qkv || !a->cos || !a->sin || !a->pos || !a->k_cache || !a->v_cache || !a->out) return B200_ERR_UNSUPPORTED;
  if (a->batch != 1 || a->n_head <= 0 || a->n_head != a->n_groups || a->head_size != 128 || a->max_seq <= 0) return B200_ERR_UNSUPPORTED;
  if (a->rope_n_elem < 0 || a->rope_n_elem > a->head_size || a->rope_n_elem % 2) return B200_ERR_UNSUPPORTED;
  if (a->table_dtype != B200_F32 && a->table_dtype != B200_BF16) return B200_ERR_UNSUPPORTED;
  AttnParams p{};
  p.qkv = (const __nv_bfloat16*)a->qkv;
  p.cos = a->cos; p.sin = a->sin; p.table_bf16 = a->table_dtype == B200_BF16; p.table_presel = a->table_preselected; p.rope_n = a->rope_n_elem;
  p.pos = a->pos; p.pos_i64 = a->pos_is_int64;
  p.kc = (__nv_bfloat16*)a->k_cache; p.vc = (__nv_bfloat16*)a->v_cache; p.out = (__nv_bfloat16*)a->out;
  p.batch = 1; p.n_head = a->n_head; p.n_groups = a->n_groups; p.hs = a->head_size; p.max_seq = a->max_seq; p.qpk = 1;
  p.nsplit = nsplit; p.head_chunks = 1;
  p.scale_log2 = (1.0f / sqrtf((float)a->head_size)) * 1.4426950408889634f;
  *out = p;
  return B200_OK;
}
}  // namespace b200

extern "C" {

// ---- allocation-free variant (no counterpart in the reference) ------------
// Same kernel, same arguments, but the split-K scratch (partial (m, l, o) + ticket counters) lives in a file-scope
// device array instead of a caller-provided workspace: for callers that cannot call cudaMalloc (CUDA graph capture,
// tight serving loops, harnesses that forbid it).  Every GPU that loads the library holds the array until exit.
// All calls share it: calls that overlap on the GPU (different streams) race; calls on one stream are ordered.
#define B200_ATTN_NOALLOC_SCRATCH_BYTES (4u << 20)
static __device__ __align__(256) unsigned char g_attn_noalloc_scratch[B200_ATTN_NOALLOC_SCRATCH_BYTES];

size_t b200_attn_decode_noalloc_max_bytes(void) { return (size_t)B200_ATTN_NOALLOC_SCRATCH_BYTES; }

int b200_attn_decode_noalloc(const b200_attn_decode_args* a, b200_stream_t stream) {
  B200_CHECK_ARG(a != nullptr, "b200_attn_decode_noalloc: null args");
  B200_CHECK_ARG(a->batch > 0 && a->n_head > 0 && a->head_size > 0, "b200_attn_decode_noalloc: bad shape");
  B200_CHECK_ARG(a->num_splits >= 0 && a->num_splits <= 64, "b200_attn_decode_noalloc: num_splits %d out of range [0, 64]", a->num_splits);
  // the split count is chosen inside b200_attn_decode when num_splits == 0: size for the most it can choose (64)
  const size_t need = b200_attn_decode_workspace_bytes(a->batch, a->n_head, a->head_size, a->num_splits > 0 ? a->num_splits : 64);
  if (need > (size_t)B200_ATTN_NOALLOC_SCRATCH_BYTES) {
    set_error("b200_attn_decode_noalloc: batch %d x n_head %d x head_size %d x %d splits needs %zu bytes of scratch, the built-in array holds %zu "
              "(use b200_attn_decode with a caller-owned workspace)", a->batch, a->n_head, a->head_size, a->num_splits > 0 ? a->num_splits : 64,
              need, (size_t)B200_ATTN_NOALLOC_SCRATCH_BYTES);
    return B200_ERR_UNSUPPORTED;
  }
  // the array's DEVICE address, resolved once per device with that device current (in host code the symbol's name is a
  // host-side stand-in, not a device pointer)
  static void* dev_ptr[64] = {};
  static std::atomic<int> dev_ready[64] = {};
  int dev = -1;
  B200_CUDA(cudaGetDevice(&dev));
  B200_CHECK_ARG(dev >= 0 && dev < 64, "b200_attn_decode_noalloc: device ordinal %d not supported", dev);
  if (!dev_ready[dev].load(std::memory_order_acquire)) {
    void* ptr = nullptr;
    B200_CUDA(cudaGetSymbolAddress(&ptr, g_attn_noalloc_scratch));
    B200_CHECK_ARG(ptr != nullptr, "b200_attn_decode_noalloc: cudaGetSymbolAddress returned NULL");
    dev_ptr[dev] = ptr;   // every thread resolves the same address: a benign race
    dev_ready[dev].store(1, std::memory_order_release);
  }
  // the ticket counters at the start of the scratch must be zero when the kernel starts; the kernel cleans up after
  // itself, but an earlier call with another shape (or an aborted one) may have left something: clear them on the stream
  const size_t cnt_bytes = (((size_t)a->batch * a->n_head * sizeof(unsigned int)) + 255) & ~(size_t)255;
  B200_CUDA(cudaMemsetAsync(dev_ptr[dev], 0, cnt_bytes, (cudaStream_t)stream));
  b200_attn_decode_args b = *a;
  b.workspace = dev_ptr[dev];
  b.workspace_bytes = (size_t)B200_ATTN_NOALLOC_SCRATCH_BYTES;
  return b200_attn_decode(&b, stream);
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
