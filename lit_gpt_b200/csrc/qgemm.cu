// Prefill GEMM: y (M, N) = x (M, K) . dequant(W)^T on the 5th-generation tensor cores (tcgen05 + TMEM), DESIGN.md §5.4.
//
// Replaces, for T > 16 rows, bnb.nn.Linear4bit.forward's "dequantize_4bit of the WHOLE weight to bf16, then F.linear"
// (call sites lit_gpt/model.py:519,619,656,713,714,716): the bf16 weight is never materialised in HBM.
// Each CTA owns a (128 * MT tokens) x (128 output features) tile and walks K in blocks of 64:
//   * the ACTIVATION tiles (128 tokens x 64 k, bf16) are fetched by TMA (cp.async.bulk.tensor.2d over a tensor map of x,
//     SWIZZLE_128B, rows past M zero-filled by the hardware) straight into the tensor core's K-major layout;
//   * 16 PRODUCER warps read the packed 4-bit weights (8 B = 16 codes per thread per k-block) and their absmax /
//     {scale, zero}, rebuild the bf16 values with exactly the dequantise kernel's arithmetic
//     (bf16_rn(code_fp32 * absmax_fp32), resp. bf16_rn(scale * (q - zero))) and write them, together with the
//     activation tile, into shared memory in the tensor core's K-major 128-byte-swizzled layout;
//   * ONE thread issues tcgen05.mma (cta_group::1, kind::f16, bf16 x bf16 -> fp32, M = 128, N = 128, K = 16) over
//     shared-memory matrix descriptors; the fp32 accumulators live in TENSOR MEMORY (128 lanes x 128 columns per
//     M sub-tile); tcgen05.commit hands each shared-memory stage back to the producers through an mbarrier;
//   * 4 EPILOGUE warps (the first four producer warps, once K is exhausted) read the accumulators with tcgen05.ld
//     (32 lanes x 32 columns per warp and instruction), add the bias, round to bf16 and store rows of y.
// With MT = 2 one dequantised weight tile feeds 256 tokens, so the CUDA-core dequant (~500 clk per k-block) is
// hidden under the tensor-core time (2 x 256 clk).
#include <cuda.h>  // CUtensorMap (type only; the encoder is resolved at run time through the runtime, no -lcuda)

#include "common.cuh"

namespace b200 {

struct GemmParams {
  const uint8_t* packed;      // reference layout: nf4/fp4 bytes (N, K/2); int4 words (N, K/8)
  const uint32_t* scale;      // fp32 absmax (N, K/64) or half2 {scale, zero} (N, K/128)
  const __nv_bfloat16* x;     // (M, K)
  const __nv_bfloat16* bias;  // (N) or null
  __nv_bfloat16* y;           // (M, N)
  int64_t M, N, K;
  int qtype;
  int ksplit;                 // > 1 (grids that would not fill the SMs): blockIdx.z owns a range of k-blocks and writes fp32 partial sums
  float* partial;             // (ksplit, M, N) fp32, summed in split order by qgemm_splitk_reduce_kernel
};

constexpr int kGProducers = 512;            // 16 warps: 4 per scheduler, enough to hide the look-up latency
constexpr int kGThreads = kGProducers + 32; // + the MMA / TMEM-allocator warp
constexpr int kGStages = 4;
constexpr int kBN = 128, kBK = 64;
constexpr int kTileBytes = 128 * kBK * 2;   // one 128-row x 64-k bf16 operand tile = 16 KB

__device__ __forceinline__ uint32_t g_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void g_mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void g_mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void g_mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  } while (!done);
}

// Shared-memory matrix descriptor, K-major operand, 128-byte swizzle (cute::UMMA::SmemDescriptor):
//   [0,14) start address >> 4 | [16,30) leading byte offset >> 4 (unused for swizzled K-major: 1) |
//   [32,46) stride byte offset >> 4 (8 rows x 128 B = 1024 B) | [46,48) version = 1 | [61,64) layout = 2 (SWIZZLE_128B)
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
  return (uint64_t)((smem_addr & 0x3FFFFu) >> 4) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// Instruction descriptor (cute::UMMA::InstrDescriptor): D = f32, A = B = bf16, both K-major, N = 128, M = 128
constexpr uint32_t kIdesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(kBN >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);

__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(da), "l"(db), "r"(kIdesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {  // arrives on `bar` when all MMAs issued so far have completed
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

// byte offset of the 16-byte chunk `c` (0..7) of row `r` inside a 128-row K-major SWIZZLE_128B tile
__device__ __forceinline__ uint32_t sw128(int r, int c) { return (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + ((c ^ (r & 7)) << 4)); }

template <int MT, bool INT4>
__global__ void __launch_bounds__(kGThreads, 1) qgemm_tcgen05_kernel(const GemmParams p, const __grid_constant__ CUtensorMap tmap_x) {
  extern __shared__ __align__(1024) uint8_t g_smem[];
  constexpr int STAGE = (MT + 1) * kTileBytes;      // MT activation tiles + 1 weight tile
  constexpr int TMEM_COLS = MT <= 2 ? 128 * MT : 512;   // a power of two (MT = 3: 384 accumulator columns of 512)
  constexpr int NSTAGE = MT <= 2 ? kGStages : 3;          // MT = 3: 64 KB stages, three of them fit
  __shared__ __align__(8) uint64_t s_bars[2 * kGStages + 1];
  __shared__ uint32_t s_tmem;
  __shared__ float s_code[16 * 32];                 // fp32 code book, one copy per lane (conflict-free look-ups)

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  uint8_t* tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(g_smem) + 1023) & ~(uintptr_t)1023);
  const uint32_t tiles_a = g_smem_u32(tiles);
  const uint32_t full0 = g_smem_u32(&s_bars[0]), empty0 = g_smem_u32(&s_bars[kGStages]), done_bar = g_smem_u32(&s_bars[2 * kGStages]);
  const int64_t n0 = (int64_t)blockIdx.x * kBN, m0 = (int64_t)blockIdx.y * (128 * MT);
  // split-K (short prompts: the output tiles alone would leave most SMs idle): this CTA's k-blocks are [kb0, kb0 + nkb)
  const int nkb_all = (int)(p.K / kBK);
  const int kb0 = (int)(((int64_t)nkb_all * blockIdx.z) / p.ksplit);
  const int nkb = (int)(((int64_t)nkb_all * (blockIdx.z + 1)) / p.ksplit) - kb0;
  constexpr bool int4 = INT4;

  if (tid == 0) {
    for (int s = 0; s < NSTAGE; ++s) { g_mbar_init(full0 + s * 8, kGProducers / 32 + 1); g_mbar_init(empty0 + s * 8, 1); }
    g_mbar_init(done_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (!int4) {
    for (int i = tid; i < 16 * 32; i += kGThreads) s_code[i] = code_value(p.qtype, i >> 5);
  }
  if (warp == kGProducers / 32) {  // the MMA warp also owns the tensor-memory allocation
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(g_smem_u32(&s_tmem)), "n"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = s_tmem;
  griddep_launch();
  griddep_wait();

  if (warp < kGProducers / 32) {
    // =========================================== PRODUCERS ======================================================
    // The activation tiles of k-block kb are requested by ONE thread with TMA (no LSU work, no registers); the packed
    // weights and scales are prefetched into registers kWLag k-blocks ahead.  full[s] collects one arrival per warp
    // (dequantised weight tile written) plus the TMA thread's expect_tx arrival; the TMA bytes complete the phase.
    constexpr int kWLag = 2;
    // weight tile: thread -> (row r = tid / 4, quarter h = tid % 4): 8 packed bytes = k [16h, 16h + 16) of the k-block
    const int wr = tid >> 2, wh = tid & 3;
    const int64_t wn = n0 + wr;
    const bool wvalid = wn < p.N;
    const uint8_t* wsrc = p.packed + (wvalid ? wn : 0) * (p.K >> 1) + wh * 8;
    const uint32_t* ssrc = p.scale + (wvalid ? wn : 0) * (int4 ? (p.K >> 7) : (p.K >> 6));
    uint2 qn[kWLag];
    uint32_t scn[kWLag];
#pragma unroll
    for (int i = 0; i < kWLag; ++i) {
      qn[i] = make_uint2(0u, 0u);
      scn[i] = 0u;
      if (i < nkb) { qn[i] = ldg_stream_v2(wsrc + (int64_t)(kb0 + i) * 32); scn[i] = ssrc[int4 ? ((kb0 + i) >> 1) : (kb0 + i)]; }
    }
    const uint32_t wdst0 = (uint32_t)(MT * kTileBytes) + sw128(wr, wh * 2), wdst1 = (uint32_t)(MT * kTileBytes) + sw128(wr, wh * 2 + 1);
    // running state instead of per-iteration index arithmetic (the loop is bound by its instruction count: 139 -> ~105 per k-block):
    // ring slot / parity, the packed-weight pointer and the k-block index of the look-ahead load (rows past N read row 0: in bounds)
    const uint8_t* wp = wsrc + (int64_t)(kb0 + kWLag) * 32;
    int kq = kb0 + kWLag;                 // absolute k-block of the next look-ahead load
    int s = 0;
    uint32_t eph = 1;                     // parity of the EMPTY phase to wait for (first lap: passes at once)
    uint32_t st_a = tiles_a;              // shared-window address of the current stage
    for (int kb = 0; kb < nkb; ++kb) {
      g_mbar_wait(empty0 + s * 8, eph);
      // ---- activation tiles: MT boxes of 128 rows x 64 k by TMA, issued by one thread ------------------------------
      if (tid == 0) {
        const uint32_t bar = full0 + s * 8;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(MT * kTileBytes) : "memory");
#pragma unroll
        for (int mt = 0; mt < MT; ++mt)
          asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                       ::"r"(st_a + (uint32_t)(mt * kTileBytes)), "l"(&tmap_x), "r"((kb0 + kb) * kBK), "r"((int)(m0 + mt * 128)), "r"(bar)
                       : "memory");
      }
      // ---- weight tile: dequantise 16 codes -> 2 chunks of 8 bf16 ----------------------------------------------------
      const uint2 q = qn[0];
      const uint32_t sc = scn[0];
#pragma unroll
      for (int i = 0; i + 1 < kWLag; ++i) { qn[i] = qn[i + 1]; scn[i] = scn[i + 1]; }
      if (kb + kWLag < nkb) { qn[kWLag - 1] = ldg_stream_v2(wp); scn[kWLag - 1] = ssrc[int4 ? (kq >> 1) : kq]; }
      wp += 32; ++kq;
      const uint32_t qw[2] = {q.x, q.y};
      if (int4) {
        // bf16_rn(scale * (q - zero)) with three instructions per code: the nibble is moved to the top four mantissa bits of 16.0f
        // (SHF + LOP3: f = 16 + q exactly, five significant bits), then ONE FFMA scale * f - scale * (16 + zero): both products are
        // exact in fp32 (11-bit fp16 scale x 6-bit integer), so the result is the exact scale * (q - zero), as before -- without the
        // quarter-rate I2F per code that bound this loop
        const __half2 h2 = *reinterpret_cast<const __half2*>(&sc);
        const float scale = __low2float(h2);
        const float c0 = -scale * (16.0f + __high2float(h2));
        auto deq = [&](uint32_t w, int e) -> float {   // element e (bits 4e .. 4e + 3) of word w
          const uint32_t sh = e <= 4 ? w << (19 - 4 * e) : w >> (4 * e - 19);
          return fmaf(scale, __uint_as_float((sh & 0x00780000u) | 0x41800000u), c0);
        };
#pragma unroll
        for (int j = 0; j < 2; ++j) {  // word j = k [16h + 8j, +8), element e in bits 4e
          uint32_t o[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) o[e] = pack_bf16(deq(qw[j], 2 * e), deq(qw[j], 2 * e + 1));
          asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(st_a + (j ? wdst1 : wdst0)), "r"(o[0]), "r"(o[1]), "r"(o[2]), "r"(o[3]) : "memory");
        }
      } else {
        // bf16_rn(code_fp32 * absmax_fp32): look-up offset = nibble << 7 | lane << 2 from the table's (link-time constant) address:
        // SHF + LOP3 + LDS per code
        const float am = __uint_as_float(sc);
        const uint32_t lane4 = (uint32_t)lane * 4u;
        auto deq = [&](uint32_t w, int bit) -> float {   // the nibble at bits bit .. bit + 3 of word w
          const uint32_t sh = bit <= 7 ? w << (7 - bit) : w >> (bit - 7);
          return *reinterpret_cast<const float*>(reinterpret_cast<const char*>(s_code) + ((sh & 0x780u) | lane4)) * am;
        };
#pragma unroll
        for (int j = 0; j < 2; ++j) {  // word j = bytes 4j..4j+3 = k [16h + 8j, +8); even element in the HIGH nibble
          uint32_t o[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) o[e] = pack_bf16(deq(qw[j], 8 * e + 4), deq(qw[j], 8 * e));
          asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(st_a + (j ? wdst1 : wdst0)), "r"(o[0]), "r"(o[1]), "r"(o[2]), "r"(o[3]) : "memory");
        }
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");       // generic-proxy writes -> visible to the tensor core
      __syncwarp();
      if (lane == 0) g_mbar_arrive(full0 + s * 8);                       // weight tile of kb
      st_a += STAGE;
      if (++s == NSTAGE) { s = 0; eph ^= 1u; st_a = tiles_a; }
    }
  } else if (lane == 0) {
    // =========================================== MMA ISSUER =====================================================
    int s = 0;
    uint32_t fph = 0;
    for (int kb = 0; kb < nkb; ++kb) {
      g_mbar_wait(full0 + s * 8, fph);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t sa = tiles_a + s * STAGE;
      const uint64_t db = umma_desc_sw128(sa + MT * kTileBytes);
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) {
        const uint64_t da = umma_desc_sw128(sa + mt * kTileBytes);
#pragma unroll
        for (int k = 0; k < kBK / 16; ++k)  // +32 B along K inside the swizzle atom = +2 in the address field
          umma_bf16(tmem + mt * 128, da + 2 * k, db + 2 * k, (kb | k) != 0);
      }
      umma_commit(empty0 + s * 8);   // stage reusable once these MMAs have read it (implies fence::before_thread_sync)
      if (++s == NSTAGE) { s = 0; fph ^= 1u; }
    }
    umma_commit(done_bar);           // accumulators complete
  }

  // =============================================== EPILOGUE =======================================================
  if (warp < 4) {
    g_mbar_wait(done_bar, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll 1
    for (int mt = 0; mt < MT; ++mt) {
      const int64_t m = m0 + mt * 128 + warp * 32 + lane;        // this thread's output row (TMEM lane = row)
#pragma unroll 1
      for (int c0 = 0; c0 < kBN; c0 += 32) {
        uint32_t v[32];
        const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)(mt * 128 + c0);
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,"
            "%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
            : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
              "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
              "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
              "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
            : "r"(taddr));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        if (m < p.M && p.ksplit > 1) {
          float* dst = p.partial + ((int64_t)blockIdx.z * p.M + m) * p.N;   // raw fp32 sums of this k range; bias and rounding in the reduce kernel
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const int64_t n = n0 + c0 + j;
            if (n < p.N) *reinterpret_cast<uint4*>(dst + n) = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
          }
        } else if (m < p.M) {
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            const int64_t n = n0 + c0 + j;
            if (n < p.N) {  // N % 8 == 0 (host check): a chunk is either fully inside or fully outside
              uint32_t o[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                float a = __uint_as_float(v[j + 2 * e]), b = __uint_as_float(v[j + 2 * e + 1]);
                if (p.bias) { a += __bfloat162float(p.bias[n + 2 * e]); b += __bfloat162float(p.bias[n + 2 * e + 1]); }
                o[e] = pack_bf16(a, b);
              }
              *reinterpret_cast<uint4*>(p.y + m * p.N + n) = make_uint4(o[0], o[1], o[2], o[3]);
            }
          }
        }
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == kGProducers / 32) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(TMEM_COLS) : "memory");
  }
}

// y = bf16(sum over the k ranges, in range order, + bias): eight outputs per thread (N % 8 == 0)
__global__ void __launch_bounds__(256) qgemm_splitk_reduce_kernel(const float* __restrict__ partial, const __nv_bfloat16* __restrict__ bias,
                                                                  __nv_bfloat16* __restrict__ y, int64_t MN, int64_t N, int ksplit) {
  griddep_launch();
  griddep_wait();
  const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (i >= MN) return;
  float acc[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[e] = 0.f;
  for (int z = 0; z < ksplit; ++z) {
    const float4 a = *reinterpret_cast<const float4*>(partial + (int64_t)z * MN + i);
    const float4 b = *reinterpret_cast<const float4*>(partial + (int64_t)z * MN + i + 4);
    acc[0] += a.x; acc[1] += a.y; acc[2] += a.z; acc[3] += a.w; acc[4] += b.x; acc[5] += b.y; acc[6] += b.z; acc[7] += b.w;
  }
  const int64_t n = i % N;
  uint32_t o[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    float a = acc[2 * e], b = acc[2 * e + 1];
    if (bias) { a += __bfloat162float(bias[n + 2 * e]); b += __bfloat162float(bias[n + 2 * e + 1]); }
    o[e] = pack_bf16(a, b);
  }
  *reinterpret_cast<uint4*>(y + i) = make_uint4(o[0], o[1], o[2], o[3]);
}

static std::atomic<int> g_gemm_mt3{1};   // development switch (b200_qgemm_tune)

// k ranges for prompts whose (M tiles) x (N / 128) CTAs leave SMs idle (up to ~1000 tokens at N = 4096): enough ranges to fill ONE wave
// (a second, partial wave costs more than the idle SMs of the first), at least one lap of the four-stage ring per range, and at
// most 64 MB of fp32 partial sums
static int gemm_ksplit(int64_t M, int64_t N, int64_t K) {
  const int nkb = (int)(K / kBK);
  const int64_t mtile = M > 512 && g_gemm_mt3.load() ? 384 : M > 128 ? 256 : 128;
  const int64_t ctas = ((N + kBN - 1) / kBN) * ((M + mtile - 1) / mtile);
  int64_t s = sm_count() / (ctas > 0 ? ctas : 1);
  if (s > nkb / kGStages) s = nkb / kGStages;
  if (s > 16) s = 16;
  while (s > 1 && s * M * N * (int64_t)sizeof(float) > (64ll << 20)) --s;
  return s < 1 ? 1 : (int)s;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) p = nullptr;
    return reinterpret_cast<EncodeTiledFn>(p);
  }();
  return fn;
}

template <int MT, bool INT4>
static cudaError_t launch_gemm_q(const GemmParams& p, cudaStream_t st) {
  constexpr int smem = (MT <= 2 ? kGStages : 3) * (MT + 1) * kTileBytes + 1024;
  static PerDeviceOnce once;
  if (once.need()) {
    cudaError_t e = cudaFuncSetAttribute(qgemm_tcgen05_kernel<MT, INT4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    once.done();
  }
  // tensor map of x (M, K) bf16 row-major: dim0 = K (contiguous), dim1 = M; box 64 x 128, 128-byte swizzle, zero fill
  EncodeTiledFn enc = encode_tiled();
  if (!enc) return cudaErrorNotSupported;
  alignas(64) CUtensorMap tmap;
  const cuuint64_t gdim[2] = {(cuuint64_t)p.K, (cuuint64_t)p.M};
  const cuuint64_t gstride[1] = {(cuuint64_t)p.K * 2};
  const cuuint32_t box[2] = {(cuuint32_t)kBK, 128u};
  const cuuint32_t estr[2] = {1u, 1u};
  if (enc(&tmap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<__nv_bfloat16*>(p.x), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
    return cudaErrorInvalidValue;
  dim3 grid((unsigned)((p.N + kBN - 1) / kBN), (unsigned)((p.M + 128 * MT - 1) / (128 * MT)), (unsigned)p.ksplit);
  cudaError_t e = launch(qgemm_tcgen05_kernel<MT, INT4>, grid, dim3(kGThreads), (size_t)smem, st, p, tmap);
  if (e != cudaSuccess || p.ksplit == 1) return e;
  const int64_t MN = p.M * p.N;
  return launch(qgemm_splitk_reduce_kernel, dim3((unsigned)((MN / 8 + 255) / 256)), dim3(256), (size_t)0, st, (const float*)p.partial, p.bias, p.y, MN, p.N, p.ksplit);
}

template <int MT>
static cudaError_t launch_gemm(const GemmParams& p, cudaStream_t st) {
  return p.qtype == B200_Q_INT4 ? launch_gemm_q<MT, true>(p, st) : launch_gemm_q<MT, false>(p, st);
}

}  // namespace b200

using namespace b200;

extern "C" {

int b200_qgemm_tune(int mt3) { g_gemm_mt3 = mt3 ? 1 : 0; return B200_OK; }

size_t b200_qgemm_workspace_bytes(int64_t M, int64_t N, int64_t K) {
  // accumulators live in tensor memory; only a launch split over k ranges (short prompts: too few output tiles for the SMs) stages fp32 partial sums
  if (M <= 0 || N <= 0 || K < kBK) return 0;
  const int s = gemm_ksplit(M, N, K);
  return s > 1 ? (size_t)s * (size_t)M * (size_t)N * sizeof(float) : 0;
}

int b200_qgemm(const b200_qgemm_args* a, b200_stream_t stream) {
  B200_CHECK_ARG(a != nullptr, "b200_qgemm: null args");
  B200_CHECK_ARG(a->packed && a->scale && a->x && a->y, "b200_qgemm: null packed/scale/x/y pointer");
  B200_CHECK_ARG(a->M > 0 && a->N > 0 && a->K > 0, "b200_qgemm: M, N, K must be positive");
  B200_CHECK_ARG(a->qtype == B200_Q_NF4 || a->qtype == B200_Q_FP4 || a->qtype == B200_Q_INT4, "b200_qgemm: bad qtype %d", a->qtype);
  const bool int4 = a->qtype == B200_Q_INT4;
  if (a->blocksize != (int4 ? 128 : 64) || a->K % (int4 ? 128 : 64) != 0 || a->N % 8 != 0) {
    set_error("b200_qgemm: needs blocksize %d, K %% %d == 0 and N %% 8 == 0 (blocksize=%d N=%lld K=%lld)", int4 ? 128 : 64, int4 ? 128 : 64,
              a->blocksize, (long long)a->N, (long long)a->K);
    return B200_ERR_UNSUPPORTED;
  }
  B200_CHECK_ARG(a->N <= 65535LL * kBN && a->M <= 65535LL * 128, "b200_qgemm: shape too large for the grid");
  B200_CHECK_ARG(((reinterpret_cast<uintptr_t>(a->packed) | reinterpret_cast<uintptr_t>(a->x) | reinterpret_cast<uintptr_t>(a->y)) & 15) == 0,
                 "b200_qgemm: packed / x / y must be 16-byte aligned");
  GemmParams p{};
  p.packed = reinterpret_cast<const uint8_t*>(a->packed);
  p.scale = reinterpret_cast<const uint32_t*>(a->scale);
  p.x = reinterpret_cast<const __nv_bfloat16*>(a->x);
  p.bias = reinterpret_cast<const __nv_bfloat16*>(a->bias);
  p.y = reinterpret_cast<__nv_bfloat16*>(a->y);
  p.M = a->M; p.N = a->N; p.K = a->K; p.qtype = a->qtype;
  // split-K only when the caller provided the workspace b200_qgemm_workspace_bytes asks for (none: one k range, as before)
  p.ksplit = gemm_ksplit(a->M, a->N, a->K);
  if (p.ksplit > 1 && (a->workspace == nullptr || a->workspace_bytes < (size_t)p.ksplit * a->M * a->N * sizeof(float) ||
                       (reinterpret_cast<uintptr_t>(a->workspace) & 15)))
    p.ksplit = 1;
  p.partial = reinterpret_cast<float*>(a->workspace);
  cudaStream_t st = (cudaStream_t)stream;
  // tokens per dequantised weight tile: 128, 256, or 384 for long prompts (the CUDA-core dequantisation is the bound: §5.4)
  const cudaError_t e = a->M > 512 && g_gemm_mt3.load() ? launch_gemm<3>(p, st) : a->M > 128 ? launch_gemm<2>(p, st) : launch_gemm<1>(p, st);
  if (e != cudaSuccess) {
    set_error("b200_qgemm: launch failed: %s", cudaGetErrorString(e));
    return B200_ERR_CUDA;
  }
  return B200_OK;
}

}  // extern "C"
