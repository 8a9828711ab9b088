// Micro-benchmark (development tool): how fast can ONE 544-thread CTA per SM pull a large buffer out of HBM into
// shared memory with (0) cp.async.bulk copies issued by one thread into an mbarrier ring, (1) cp.async 16 B by
// every thread (commit groups), (2) plain LDG.128 into registers?  Consumers never read the data: this is the
// delivery ceiling of each mechanism.   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o stream_micro stream_micro.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(b), "r"(c)); }
__device__ __forceinline__ void mbar_expect(uint32_t b, uint32_t n) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(n) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint32_t b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(b) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t b, uint32_t par) {
  uint32_t d;
  do { asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(d) : "r"(b), "r"(par) : "memory"); } while (!d);
}
__device__ __forceinline__ void bulk(uint32_t dst, const void* src, uint32_t n, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(n), "r"(bar) : "memory");
}

// mode 0: bulk ring. stage bytes = sb, copies per stage = cps, stages = S, issuing lanes = nl (lane i issues copies i, i+nl, ...)
__global__ void __launch_bounds__(544, 1) k_bulk(const uint8_t* src, int64_t per_cta, int sb, int cps, int S, int nl, unsigned* sink) {
  extern __shared__ __align__(128) uint8_t sm[];
  __shared__ __align__(8) uint64_t bars[32];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t full = s32(&bars[0]), empty = s32(&bars[16]);
  if (tid == 0) { for (int s = 0; s < S; ++s) { mbar_init(full + 8 * s, 1); mbar_init(empty + 8 * s, 16); } asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  __syncthreads();
  const uint8_t* base = src + (int64_t)blockIdx.x * per_cta;
  const int nst = (int)(per_cta / sb);
  if (warp == 16) {
    const int cb = sb / cps;
    for (int n = 0; n < nst; ++n) {
      const int s = n % S;
      if (n >= S) { if (lane == 0) mbar_wait(empty + 8 * s, ((n / S) - 1) & 1); __syncwarp(); }
      if (lane == 0) mbar_expect(full + 8 * s, sb);
      __syncwarp();
      for (int c = lane; c < cps; c += nl) if (lane < nl) bulk(s32(sm) + s * sb + c * cb, base + (int64_t)n * sb + c * cb, cb, full + 8 * s);
    }
  } else {
    unsigned acc = 0;
    for (int n = 0; n < nst; ++n) {
      const int s = n % S;
      mbar_wait(full + 8 * s, (n / S) & 1);
      acc += sm[s * sb + tid * 4];
      __syncwarp();
      if (lane == 0) mbar_arrive(empty + 8 * s);
    }
    if (acc == 0xdeadbeef) sink[0] = acc;
  }
}

// mode 1: cp.async 16 B per thread, G groups in flight, each group = 512 threads * 16 B * U
template <int U>
__global__ void __launch_bounds__(512, 1) k_cpasync(const uint8_t* src, int64_t per_cta, int G, unsigned* sink) {
  extern __shared__ __align__(128) uint8_t sm[];
  const int tid = threadIdx.x;
  const uint8_t* base = src + (int64_t)blockIdx.x * per_cta;
  const int gb = 512 * 16 * U;
  const int ng = (int)(per_cta / gb);
  for (int g = 0; g < ng + G; ++g) {
    if (g < ng) {
#pragma unroll
      for (int u = 0; u < U; ++u)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s32(sm) + (g % G) * gb + (u * 512 + tid) * 16), "l"(base + (int64_t)g * gb + (u * 512 + tid) * 16) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group %0;" ::"n"(7) : "memory");  // G <= 8 groups in flight
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  if (sm[tid] == 0xab && per_cta == 1) sink[0] = 1;
}

// mode 2: LDG.128, U independent loads per thread per iteration
template <int U>
__global__ void __launch_bounds__(512, 1) k_ldg(const uint8_t* src, int64_t per_cta, unsigned* sink) {
  const int tid = threadIdx.x;
  const uint4* base = reinterpret_cast<const uint4*>(src + (int64_t)blockIdx.x * per_cta);
  const int64_t n16 = per_cta / 16;
  unsigned acc = 0;
  for (int64_t i = tid; i + (U - 1) * 512 < n16; i += 512 * U) {
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(base + i + u * 512));
#pragma unroll
    for (int u = 0; u < U; ++u) acc += v[u].x ^ v[u].w;
  }
  if (acc == 0xdeadbeef) sink[0] = acc;
}

int main() {
  const int sms = 148;
  const int64_t per_cta = 8 << 20;           // 8 MB per CTA -> 1.18 GB total: far larger than L2
  const int64_t total = per_cta * sms;
  uint8_t* buf; unsigned* sink;
  cudaMalloc(&buf, total); cudaMalloc(&sink, 4); cudaMemset(buf, 1, total);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  auto report = [&](const char* name, float ms) { printf("%-58s %8.1f us  %7.1f GB/s\n", name, ms * 1e3, total / ms / 1e6); fflush(stdout); };
  const int smem = 200 * 1024;
  cudaFuncSetAttribute(k_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  struct Cfg { int sb, cps, S, nl; } cfgs[] = {{18432, 2, 6, 1}, {18432, 2, 10, 1}, {18432, 8, 10, 1}, {18432, 8, 10, 8}, {18432, 32, 10, 32}, {9216, 1, 20, 1}, {36864, 2, 5, 1}, {36864, 16, 5, 16}, {4608, 1, 40, 1}};
  for (auto c : cfgs) {
    if (c.S > 16) c.S = 16;
    const int64_t pc = per_cta / c.sb * c.sb;
    k_bulk<<<sms, 544, c.S * c.sb>>>(buf, pc, c.sb, c.cps, c.S, c.nl, sink);
    cudaEventRecord(e0); k_bulk<<<sms, 544, c.S * c.sb>>>(buf, pc, c.sb, c.cps, c.S, c.nl, sink); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    char name[128]; snprintf(name, sizeof name, "bulk ring: stage %5d B, %2d copies/stage, %2d stages, %2d lanes", c.sb, c.cps, c.S, c.nl);
    report(name, ms * (float)per_cta / pc);
  }
  cudaFuncSetAttribute(k_cpasync<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(k_cpasync<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  for (int G : {2, 4, 6}) {
    float ms;
    k_cpasync<4><<<sms, 512, G * 512 * 16 * 4>>>(buf, per_cta, G, sink);
    cudaEventRecord(e0); k_cpasync<4><<<sms, 512, G * 512 * 16 * 4>>>(buf, per_cta, G, sink); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    char name[128]; snprintf(name, sizeof name, "cp.async 16 B: 32 KB groups, ring of %d (8 groups in flight)", G);
    report(name, ms);
  }
  {
    float ms;
    k_ldg<4><<<sms, 512>>>(buf, per_cta, sink);
    cudaEventRecord(e0); k_ldg<4><<<sms, 512>>>(buf, per_cta, sink); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1); report("LDG.128 x4 per thread, 512 threads/SM", ms);
    k_ldg<8><<<sms, 512>>>(buf, per_cta, sink);
    cudaEventRecord(e0); k_ldg<8><<<sms, 512>>>(buf, per_cta, sink); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1); report("LDG.128 x8 per thread, 512 threads/SM", ms);
  }
  cudaError_t e = cudaDeviceSynchronize();
  printf("status: %s\n", cudaGetErrorString(e));
  return 0;
}
