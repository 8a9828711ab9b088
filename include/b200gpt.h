/* b200gpt — C ABI of the B200-native (sm_100a) quantized-decode hot path of lit-gpt.
 *
 * Every entry point takes raw DEVICE pointers, sizes and a CUDA stream (`cudaStream_t` passed as void*),
 * launches asynchronously on that stream, never allocates, never synchronises, and returns 0 on success.
 * On failure it returns a B200_ERR_* code and `b200_last_error()` (thread-local) holds the message.
 * Caller-owned workspaces are sized by the *_workspace_bytes helpers.
 *
 * Each function cites the reference interface it replaces (file:line into yuanyaaa/lit-gpt).  The reference's
 * own quantized arithmetic lives in bitsandbytes==0.41.0 (requirements-all.txt:3) behind
 * lightning.fabric.plugins.BitsandbytesPrecision (generate/base.py:129-134); see INTEGRATION.md for the
 * binding a maintainer would add.
 */
#ifndef B200GPT_H
#define B200GPT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200GPT_ABI_VERSION 1

typedef void* b200_stream_t; /* cudaStream_t */

enum b200_status { B200_OK = 0, B200_ERR_INVALID = 1, B200_ERR_CUDA = 2, B200_ERR_UNSUPPORTED = 3 };
enum b200_dtype { B200_BF16 = 0, B200_F16 = 1, B200_F32 = 2 };
/* 4-bit weight formats.  NF4/FP4: bitsandbytes block-64 codebooks; INT4: this project's group-128 format
 * (the reference has no gptq/int4 code). */
enum b200_qtype { B200_Q_NF4 = 0, B200_Q_FP4 = 1, B200_Q_INT4 = 2 };
/* GEMV epilogues */
enum b200_epilogue { B200_EPI_STORE = 0, B200_EPI_RESIDUAL = 1, B200_EPI_SWIGLU = 2 };

/* ---- library ------------------------------------------------------------------------------------------ */
int b200_abi_version(void);
const char* b200_last_error(void);
/* fills SM count and compute capability of the current device */
int b200_device_info(int* sm_count, int* cc_major, int* cc_minor);
/* enable/disable programmatic dependent launch (griddepcontrol) on every kernel of the library; returns old value */
int b200_set_pdl(int enabled);
/* number of kernels launched by this library since load (the `gpu_launches` evidence of bench.py) */
int64_t b200_launch_count(void);
/* tuning override for the fast GEMV: ksteps in {0(auto),1,2,4}, ctas_per_sm in {0(auto),1,2,3,4}; returns 0 */
int b200_qgemv_tune(int ksteps, int ctas_per_sm, int force_generic);

/* ---- quantisation (replaces bitsandbytes.functional.quantize_4bit / dequantize_4bit reached through
 *      Params4bit.cuda(); reference call sites generate/base.py:168, generate/tp.py:190, lit_gpt/lora.py:156-160) */
/* w: `numel` elements of w_dtype, flattened row-major.  packed: (numel+1)/2 bytes, even element in the HIGH
 * nibble.  absmax: ceil(numel/blocksize) floats.  qtype NF4|FP4.  blocksize must be even (reference uses 64). */
int b200_quantize_4bit(const void* w, int w_dtype, int64_t numel, int blocksize, int qtype, uint8_t* packed,
                       float* absmax, b200_stream_t stream);
int b200_dequantize_4bit(const uint8_t* packed, const float* absmax, int64_t numel, int blocksize, int qtype,
                         void* out, int out_dtype, b200_stream_t stream);
/* Double quantisation of an absmax vector (--quantize bnb.nf4-dq / bnb.fp4-dq; bitsandbytes quantize_4bit with
 * compress_statistics=True, reference call sites generate/base.py:105,129-134): blocks of 256, `offset` (the mean,
 * computed by the caller) is subtracted, block absmax -> absmax2, 8-bit index into the 256-entry dynamic code
 * (code256, sorted ascending) -> q_out, and absmax_deq_out = code[q] * absmax2 + offset, the value every later
 * dequantise / GEMV of the layer uses.  n values in, ceil(n / 256) absmax2 out. */
int b200_double_quant_absmax(const float* absmax, int64_t n, const float* code256, float offset, uint8_t* q_out, float* absmax2_out,
                             float* absmax_deq_out, b200_stream_t stream);

/* int4 group-128 (asymmetric RTN, GPTQ storage order): qweight uint32 (N, ceil(K/8)), scales fp16 (N, G),
 * qzeros uint32 (N, ceil(G/8)), G = ceil(K/group).  `sz` (optional, may be NULL) receives the kernel-side
 * interleaved half2 {scale, zero} per (row, group) that b200_qgemv consumes. */
int b200_quantize_int4(const void* w, int w_dtype, int64_t N, int64_t K, int group, uint32_t* qweight,
                       uint16_t* scales, uint32_t* qzeros, uint32_t* sz, b200_stream_t stream);
int b200_dequantize_int4(const uint32_t* qweight, const uint16_t* scales, const uint32_t* qzeros, int64_t N,
                         int64_t K, int group, void* out, int out_dtype, b200_stream_t stream);

/* ---- fused dequant-GEMV, batch 1..16 (replaces bnb.nn.Linear4bit.forward -> kgemm_4bit_inference_naive /
 *      dequantize+F.linear at lit_gpt/model.py:619,656,713,714,716,519) ------------------------------------
 * y[b, n] = epilogue( sum_k xin[b, k] * W[n, k] ),  xin = prologue(x):
 *   prologue: xin = x (+ delta if delta != NULL, rounded to bf16); if x_out != NULL the summed rows are written
 *             there by one CTA (x_out must not alias x); if norm_weight != NULL xin = RMSNorm(xin) with
 *             lit_gpt/rmsnorm.py:19-25 rounding (fp32 math, one cast).
 *   W:        qtype NF4/FP4: `packed` + `absmax` (blocksize 64) as produced by b200_quantize_4bit on the
 *             row-major (N, K) weight; INT4: qweight + sz.  For B200_EPI_SWIGLU a second matrix (packed2/scale2)
 *             of the same shape is given and y[b,n] = silu(W·xin) * (W2·xin) with the reference's bf16 rounding
 *             points (lit_gpt/model.py:712-716).
 *   epilogue: STORE: y = acc (+ bias); RESIDUAL: y = bf16(acc + bias) + residual (lit_gpt/model.py:591-592).
 * x, delta, x_out, residual, y, bias, norm_weight are bf16.  batch <= 16.  workspace: b200_qgemv_workspace_bytes().
 */
typedef struct b200_qgemv_args {
  const void* packed;     /* uint8 (NF4/FP4) or uint32 (INT4) */
  const void* scale;      /* float absmax (NF4/FP4) or half2 sz (INT4) */
  const void* packed2;    /* second matrix for SWIGLU, else NULL */
  const void* scale2;
  int qtype;
  int blocksize;          /* 64 for NF4/FP4, 128 for INT4 */
  int64_t N, K;
  const void* x;          /* (batch, K) bf16 */
  const void* delta;      /* optional (batch, K) bf16 */
  void* x_out;            /* optional (batch, K) bf16 */
  const void* norm_weight;/* optional (K) bf16 */
  float norm_eps;
  int batch;
  int epilogue;
  const void* bias;       /* optional (N) bf16 */
  const void* residual;   /* (batch, N) bf16 for EPI_RESIDUAL; may alias y */
  void* y;                /* (batch, N) bf16 */
  void* workspace;
  size_t workspace_bytes;
} b200_qgemv_args;
size_t b200_qgemv_workspace_bytes(void);
int b200_qgemv(const b200_qgemv_args* args, b200_stream_t stream);

/* ---- tiled weight layout + streaming GEMV (the decode hot path; csrc/tile.cu, csrc/qgemv_stream.cu) -----------
 * b200_tile_4bit permutes the reference-layout arrays (packed + absmax of b200_quantize_4bit, or qweight + sz of
 * b200_quantize_int4) into per-(16-row tile, 64-wide k-block) records laid out in warp-consumption order;
 * b200_untile_4bit is the exact inverse (so the bitsandbytes-layout bytes are always recoverable bit for bit).
 * K must be a multiple of 64 (NF4/FP4, blocksize 64) or 128 (INT4, group 128).  b200_tiled_bytes returns 0 for
 * shapes the layout does not cover. */
size_t b200_tiled_bytes(int64_t N, int64_t K, int qtype);
int b200_tile_4bit(const void* packed, const void* scale, int64_t N, int64_t K, int qtype, int blocksize, void* tiled,
                   b200_stream_t stream);
int b200_untile_4bit(const void* tiled, int64_t N, int64_t K, int qtype, int blocksize, void* packed, void* scale,
                     b200_stream_t stream);
/* Same contract as b200_qgemv (prologue / epilogues / rounding points), reading tiled matrices; no workspace.
 * `prefetch`/`prefetch_bytes` (optional): a range the kernel asks the L2 to fetch (cp.async.bulk.prefetch.L2) -
 * the weights of the NEXT kernel in the decode graph - so HBM stays busy across the kernel boundary. */
/* Tensor-parallel hand-off fused into the GEMV (replaces the all_reduce hook of generate/tp.py:73-74 on the decode path;
 * DESIGN.md §6).  PRODUCER = the row-parallel Linear (attn.proj, mlp.proj): its epilogue stores this rank's partial
 * into slot `rank` of EVERY rank's buffer over NVLink peer memory as self-validating 32-bit packets
 * `tag << 16 | bf16 bits` (one store per output element, no fence, no separate flag: the "LL" scheme).
 * CONSUMER = the next Linear (whose `delta` the all-reduced tensor would have been): its prologue polls the `world`
 * local slots until every packet carries the expected tag, adds them in rank order in fp32, rounds once to bf16 and
 * carries on as with `delta`; its last CTA bumps the buffer's sequence counter.  tag = seq % 65535 + 1, where seq is
 * the buffer's LOCAL counter (all ranks run the same launch sequence, so the counters stay in lock step).
 * Consecutive hand-offs must alternate between two buffers.  Counters are uint32 and zero before first use, slots are
 * zero before first use.  Pointers come from b200_p2p_alloc / b200_p2p_open. */
typedef struct b200_tp_handoff {
  int world, rank;
  void* peer_y[8];        /* producer: slot `rank` (batch, N) uint32 packets in the buffer of rank r, own rank included */
  void* reduced;          /* consumer, optional: LOCAL buffer of >= batch * K uint32 packets.  When set the exchange is summed in two
                             stages: every CTA adds the `world` slots for ITS slice of the vector only and republishes the sums here as
                             packets with the same tag, then all CTAs read this one buffer: world x fewer L2 reads per CTA (at world 8
                             each CTA otherwise reads 8 x K x 4 bytes), one more local round trip.  Same arithmetic, bit-identical. */
  void* reserved[7];      /* must be NULL */
  void* send_seq;         /* producer: LOCAL sequence counter of the buffer written; NULL = this launch is not a producer */
  const void* slots;      /* consumer: local slots, rank r at slots + r * slot_stride bytes; NULL = not a consumer */
  int64_t slot_stride;    /* >= batch * K * 4 bytes */
  void* recv_seq;         /* consumer: LOCAL sequence counter of the buffer read (bumped by the launch's last CTA) */
  void* recv_done;        /* consumer: local CTA counter */
} b200_tp_handoff;
/* All-gather over the same slots (`lm_head` sharded by rows under tensor parallelism; the reference replicates it,
 * generate/tp.py:77-80).  The sharded Linear's GEMV is launched as a PRODUCER (tp->send_seq / peer_y set, y may be NULL): rank r's
 * (batch, N / world) outputs land in slot r of every rank's gather buffer.  This call is the CONSUMER: it polls the `world` slots of
 * the local buffer (consumer fields of `h`), unpacks them into out (batch, world * n_per_rank) bf16 in rank order -- i.e. the
 * unsharded Linear's output -- and bumps the buffer's sequence counter.  The gather buffer is a third hand-off buffer with its own
 * counters (it must not be one of the two alternating branch buffers).  error_word: optional, set to 2 on a bounded-wait time-out. */
int b200_tp_gather(const b200_tp_handoff* h, int batch, int64_t n_per_rank, void* out, void* error_word, b200_stream_t stream);
/* Symmetric peer buffers (setup time, not per token): cudaMalloc + zero + IPC export / import / release. */
int b200_p2p_alloc(size_t bytes, void** ptr, void* handle64);
int b200_p2p_open(const void* handle64, void** ptr);
int b200_p2p_close(void* ptr);
int b200_p2p_free(void* ptr);

/* Optional hint for the launch that precedes a decode-attention step (the fused QKV projection): once its own weight stream
 * has been issued, the kernel asks the L2 for the filled part of this layer's K / V cache (cp.async.bulk.prefetch.L2), so that
 * the attention kernel's dependent loads hit L2 instead of HBM.  `chunks` = batch * n_query_groups contiguous (Tmax, hs) planes
 * per tensor, `chunk_stride` bytes apart; rows [0, *pos) of `row_bytes` each are fetched.  A hint only: no effect on results. */
struct b200_attn_decode_args;
typedef struct b200_kv_hint {
  const void* k;
  const void* v;
  const void* pos;       /* device scalar, int32 or int64 */
  int pos_is_int64;
  int chunks;
  int64_t chunk_stride;
  int row_bytes;
  int max_rows;
} b200_kv_hint;

typedef struct b200_qgemv_tiled_args {
  const void* tiled;
  const void* tiled2;      /* second matrix for SWIGLU, else NULL */
  int qtype;
  int blocksize;           /* 64 for NF4/FP4, 128 for INT4 */
  int64_t N, K;
  const void* x;           /* (batch, K) bf16 */
  const void* delta;       /* optional (batch, K) bf16 */
  void* x_out;             /* optional (batch, K) bf16 */
  const void* norm_weight; /* optional (K) bf16 */
  float norm_eps;
  int batch;
  int epilogue;
  const void* bias;        /* optional (N) bf16 */
  const void* bias2;       /* optional (N) bf16, SWIGLU second matrix */
  const void* residual;    /* (batch, N) bf16 for EPI_RESIDUAL; may alias y */
  void* y;                 /* (batch, N) bf16 */
  const void* prefetch;
  int64_t prefetch_bytes;
  void* workspace;         /* zero-initialised once by the caller, >= b200_qgemv_tiled_workspace_bytes(); holds the
                              partial sums / arrival counters of row pairs cut by a stream-K boundary (self-cleaning) */
  size_t workspace_bytes;
  const struct b200_tp_handoff* tp; /* optional tensor-parallel hand-off (below); NULL on one GPU */
  const b200_kv_hint* kv_hint; /* optional (above); NULL = none */
  const struct b200_attn_decode_args* attn; /* optional: the decode-attention step that consumes this launch's output, FUSED into
                              the same kernel (replaces lit_gpt/model.py:619 + 624-653 with one launch): y is the interleaved QKV row,
                              attn->qkv must equal y, attn->out receives (1, n_head * head_size).  Batch 1, n_head == n_query_groups,
                              head size 128, STORE epilogue; each head's attention runs on the CTA of the same index as soon as the
                              head's q / k / v rows are stored (per-head counters in the workspace), four key ranges on four
                              128-thread sub-blocks merged in shared memory.  Anything else: B200_ERR_UNSUPPORTED, nothing launched
                              (call b200_attn_decode separately).  attn->workspace is not used. */
  int fold;                /* 0 / 1: plain.  F = 2, 4: `tiled` holds the matrix VIEWED as (N / F, F * K), i.e. b200_tile_4bit(N / F, F * K)
                              of the same flat nf4 / fp4 buffers: for row-parallel shards whose K is not a multiple of the
                              64-element block while F * K is (generate/tp.py:28-45 cuts mlp.proj to 11008 / 8 = 1376 columns, and
                              bitsandbytes quantizes the shard's FLAT storage, so blocks straddle rows).  N, K, x and y keep the
                              true shapes; STORE / RESIDUAL epilogues, no norm / delta / tp consumer, batch * F <= 16. */
} b200_qgemv_tiled_args;
/* 1 when b200_qgemv_tiled would run these arguments on a streaming kernel, 0 when the shape is outside every plan (nothing
 * is launched, pointers are not dereferenced): lets the caller choose the generic b200_qgemv / b200_qgemm up front. */
int b200_qgemv_tiled_supported(const b200_qgemv_tiled_args* args);
/* Bounded waits instead of hangs or traps: a kernel that waited ~2 s for a partner (a tensor-parallel peer's packets, a
 * stream-K partner CTA) sets word 0 of the workspace and carries on.  This call synchronizes `stream`, returns the word in
 * *code (0 = none, 1 = stream-K partner, 2 = tensor-parallel peer) and clears it.  generate() checks it at its sync points. */
int b200_qgemv_tiled_error(void* workspace, b200_stream_t stream, unsigned int* code);
size_t b200_qgemv_tiled_workspace_bytes(void);
int b200_qgemv_tiled(const b200_qgemv_tiled_args* args, b200_stream_t stream);
/* tuning: depth in {0(auto),2,4,6} items loaded ahead per warp; ctas_per_sm in {0(auto),1,2}; l2_prefetch 0/1 */
int b200_qgemv_tiled_tune(int depth, int ctas_per_sm, int l2_prefetch);
/* diagnostics (no counterpart in the reference): the next `slots` b200_qgemv_tiled launches record eight per-CTA phase time
 * stamps (%globaltimer, ns) into buf[slots][160][8] (uint64, device memory, caller-owned, zeroed by the caller);
 * buf = NULL switches it off.  tools/trace_step.py turns them into the per-launch timeline quoted in DESIGN.md. */
int b200_qgemv_tiled_trace(void* buf, int slots);

/* ---- RMSNorm (replaces lit_gpt/rmsnorm.py:19-25 and the residual add at lit_gpt/model.py:591) ------------
 * y = RMSNorm(x (+ delta)) * weight; if sum_out != NULL it receives bf16(x + delta).  rows x dim, bf16. */
int b200_rmsnorm(const void* x, const void* delta, const void* weight, float eps, void* y, void* sum_out,
                 int64_t rows, int64_t dim, b200_stream_t stream);

/* ---- fused RoPE + KV-cache append + split-K flash-decode attention, T = 1 (replaces lit_gpt/model.py:624-653:
 *      view/permute/split, apply_rope 767-773, KVCache.forward 788-795, scaled_dot_product_attention 658-665)
 * qkv: (batch, (n_head + 2*n_groups) * head_size) bf16 in the reference's per-group interleaved order
 *      [q * q_per_kv, k, v] x n_groups (lit_gpt/model.py:621-628).
 * cos/sin: the model's own tables (max_seq, rope_n_elem), fp32 or bf16 (never recomputed in-kernel).
 * pos: device int64/int32 scalar = position being written (input_pos); keys 0..pos are attended.
 * k_cache/v_cache: (batch, n_groups, max_seq, head_size) bf16, GQA-native (n_groups heads).
 * out: (batch, n_head * head_size) bf16.  head_size in {64, 128}. */
typedef struct b200_attn_decode_args {
  const void* qkv;
  const void* cos;
  const void* sin;
  int table_dtype;   /* B200_F32 or B200_BF16 */
  int table_preselected; /* 1: cos/sin point at the row already gathered for this token (model.py:505-506); 0: full tables indexed by pos */
  int rope_n_elem;
  const void* pos;   /* device scalar */
  int pos_is_int64;
  void* k_cache;
  void* v_cache;
  void* out;
  int batch, n_head, n_groups, head_size, max_seq;
  int num_splits;    /* 0 = auto */
  void* workspace;
  size_t workspace_bytes;
} b200_attn_decode_args;
size_t b200_attn_decode_workspace_bytes(int batch, int n_head, int head_size, int max_splits);
int b200_attn_decode(const b200_attn_decode_args* args, b200_stream_t stream);
/* development / test knob: 1 = always use the scalar kernel with the workspace merge, even for grouped-query shapes that
 * the mma.sync kernel covers (head_size 128, 2 <= q_per_kv <= 16) and for split counts <= 8 that would otherwise run as a
 * thread-block cluster merging through distributed shared memory.  Returns the previous setting. */
int b200_attn_decode_tune(int force_scalar);

/* Allocation-free variant of b200_attn_decode.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same arguments and results as b200_attn_decode, except that `workspace` / `workspace_bytes` are
 * ignored: the split-K scratch lives in a static device array of b200_attn_decode_noalloc_max_bytes() = 4 MiB that every
 * GPU loading the library holds until exit.  Shapes whose scratch need
 * b200_attn_decode_workspace_bytes(batch, n_head, head_size, num_splits ? num_splits : 64) exceeds that are rejected with
 * B200_ERR_UNSUPPORTED and a message.  No cudaMalloc / cudaFree / allocating routine is called; one cudaMemsetAsync on
 * `stream` clears the counters that must start at zero.  All calls share the array: calls that overlap on the GPU
 * (different streams, or different threads on different streams) race; calls on one stream are ordered and safe. */
size_t b200_attn_decode_noalloc_max_bytes(void);
int b200_attn_decode_noalloc(const b200_attn_decode_args* args, b200_stream_t stream);

/* ---- prefill helper: RoPE + KV-cache write for T > 1 tokens at positions pos[0..T) (lit_gpt/model.py:624-649).
 * table_preselected = 1: cos/sin are (T, rope_n_elem) rows already gathered by input_pos; 0: full tables.
 * Writes roped q to q_out (batch, n_head, T, head_size) and the roped k / v rows into the GQA-native caches. */
int b200_rope_kv_append(const void* qkv, const void* cos, const void* sin, int table_dtype, int table_preselected,
                        int rope_n_elem, const void* pos, int pos_is_int64, void* k_cache, void* v_cache, void* q_out, int batch,
                        int T, int n_head, int n_groups, int head_size, int max_seq, b200_stream_t stream);

/* ---- prefill attention (T > 1) over the GQA-native caches: replaces F.scaled_dot_product_attention at
 * lit_gpt/model.py:658-665 for the cached multi-token call (model.py:646-650).  q is (batch, n_head, T, head_size),
 * already roped (b200_rope_kv_append); query row t attends to cache positions j <= pos[t] (the mask row
 * model.py:509 selects), so the caches must already hold positions up to max(pos).  fp32 online softmax.
 * y is written as (batch, T, n_head * head_size), the layout attn.proj consumes (model.py:653).
 * head_size 64 or 128. */
int b200_attn_prefill(const void* q, const void* k_cache, const void* v_cache, const void* pos, int pos_is_int64, void* y,
                      int batch, int T, int n_head, int n_groups, int head_size, int max_seq, float scale,
                      b200_stream_t stream);

/* ---- small glue kernels so a whole decode step is one CUDA graph of this library's kernels ---------------- */
/* out[r, :] = wte[idx[r], :]  (lit_gpt/model.py:515) */
/* y = silu(a) * b over n bf16 elements (replaces `F.silu(x_fc_1) * x_fc_2`, lit_gpt/model.py:715, on the prefill path; the decode GEMV
 * has it as its SWIGLU epilogue).  Same rounding points as the reference's two ATen kernels: silu to bf16, product to bf16. */
int b200_swiglu(const void* a, const void* b, void* y, int64_t n, b200_stream_t stream);
int b200_embedding(const void* idx, int idx_is_int64, const void* wte, void* out, int64_t rows, int64_t dim,
                   int64_t vocab, b200_stream_t stream);
/* greedy sampling: token[b] = argmax(logits[b, :vocab]) (generate/base.py:41).  If pos_inout != NULL: pos += 1
 * (generate/base.py:92) and, if history != NULL, history[pos] = token[0] (the torch.cat of generate/base.py:93
 * done on the device), so a whole decode loop needs no host round trip. */
int b200_argmax(const void* logits, int64_t batch, int64_t vocab, int64_t stride, void* token_out,
                int token_is_int64, void* pos_inout, int pos_is_int64, void* history, int64_t history_len,
                b200_stream_t stream);
/* stochastic sampling of batch row 0: sample() / multinomial_num_samples_1() of generate/base.py:22-41 as one launch.
 * logits: vocab values (bf16 or fp32).  top_k <= 0 or >= vocab: no truncation, else only the top_k largest logits are
 * kept (ties at the threshold in index order).  temperature > 0: token = argmax_i softmax(l / T)_i / e_i with
 * e_i ~ Exp(1) (the draw the reference uses under torch.compile, base.py:23-26); temperature <= 0: argmax.
 * rng_state: device int64[2] = {seed, offset}; e_i comes from Philox4x32-10(seed, offset, i) and offset is advanced
 * by one per call (so a captured CUDA graph draws fresh numbers on every replay).  noise (optional, fp32[vocab]):
 * use these Exp(1) values instead (tests).  token / pos / history bookkeeping exactly as b200_argmax. */
int b200_sample(const void* logits, int logits_dtype, int64_t vocab, float temperature, int top_k, void* rng_state,
                const void* noise, void* token_out, int token_is_int64, void* pos_inout, int pos_is_int64, void* history,
                int64_t history_len, b200_stream_t stream);

/* ---- prefill GEMM: y (M, N) = x (M, K) · dequant(W)^T, tcgen05/TMEM mainloop (replaces dequantize_4bit +
 *      F.linear for T > 1 at the same call sites).  Implemented in csrc/qgemm_tcgen05.cu. */
typedef struct b200_qgemm_args {
  const void* packed;
  const void* scale;
  int qtype;
  int blocksize;
  int64_t M, N, K;
  const void* x; /* (M, K) bf16 */
  const void* bias;
  void* y;       /* (M, N) bf16 */
  void* workspace;
  size_t workspace_bytes;
} b200_qgemm_args;
size_t b200_qgemm_workspace_bytes(int64_t M, int64_t N, int64_t K);
int b200_qgemm(const b200_qgemm_args* args, b200_stream_t stream);
/* tuning: mt3 = 1 (default): prompts of more than 512 tokens share each dequantised weight tile among 384 tokens (three 128-row
 * accumulators in tensor memory, three 64 KB stages); 0: 256 tokens as in round 1 */
int b200_qgemm_tune(int mt3);

/* ---- bnb.int8: LLM.int8() inference Linear (replaces bitsandbytes.nn.Linear8bitLt(has_fp16_weights=False,
 *      threshold=6.0) installed by BitsandbytesPrecision("int8") at generate/base.py:129-134; arithmetic restated in
 *      oracle/quant_ref.py:310-345).  Implemented in csrc/int8.cu.
 *      b200_quantize_int8: CB = rint(W * 127 / absmax_row) int8 (N, K), SCB = absmax_row fp32 (N).
 *      b200_linear_int8:   y (M, N) fp16 = LLM.int8()(x (M, K) fp16): outlier columns (any |x| >= threshold) through an
 *      fp16 side product, the rest int8 x int8 -> int32 on the tensor cores.  K must be a multiple of 64
 *      (else B200_ERR_UNSUPPORTED).  The workspace holds the quantised activations; nothing is allocated inside. */
int b200_quantize_int8(const void* w, int w_dtype, int64_t N, int64_t K, int8_t* cb, float* scb, b200_stream_t stream);
int b200_dequantize_int8(const int8_t* cb, const float* scb, int64_t N, int64_t K, void* out, int out_dtype, b200_stream_t stream);
typedef struct b200_linear_int8_args {
  const void* x;      /* (M, K) fp16 */
  const void* cb;     /* (N, K) int8 */
  const void* scb;    /* (N) fp32 */
  const void* bias;   /* (N) fp16 or NULL */
  void* y;            /* (M, N) fp16 */
  int64_t M, N, K;
  float threshold;    /* 6.0 in the reference; <= 0 disables the outlier path */
  void* workspace;
  size_t workspace_bytes;
} b200_linear_int8_args;
size_t b200_linear_int8_workspace_bytes(int64_t M, int64_t K);
int b200_linear_int8(const b200_linear_int8_args* args, b200_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* B200GPT_H */
