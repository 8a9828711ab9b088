// Micro-benchmark: what does one kernel boundary of the decode chain cost?  A CUDA graph of NL dependent launches (programmatic
// dependent launch on every edge), each of which reads the previous launch's 8 KB output vector in every CTA, reduces it and
// writes its own slice: the dependency pattern of consecutive GEMVs with no weight stream at all.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/micro/launch_floor tools/micro/launch_floor.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>

__global__ void chain_kernel(const float* __restrict__ in, float* __restrict__ out, int n, int mode, int early) {
  extern __shared__ float sm[];
  if (early) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (mode == 0) return;                                  // empty kernel
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (mode == 1) { if (threadIdx.x == 0) out[blockIdx.x] = 1.f; return; }   // wait + one store
  float s = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) s += __ldcg(in + i);   // read the whole previous output (L2)
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = threadIdx.x < (blockDim.x >> 5) ? sm[threadIdx.x] : 0.f;
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    const int per = (n + gridDim.x - 1) / gridDim.x;
    for (int i = threadIdx.x; i < per; i += 32) {
      const int idx = blockIdx.x * per + i;
      if (idx < n) out[idx] = t * 1e-6f + (float)idx;
    }
  }
  if (!early) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

static float run(int grid, int threads, int smem, int mode, int early, int pdl, int NL) {
  float *a, *b;
  const int n = 2048;
  cudaMalloc(&a, n * 4); cudaMalloc(&b, n * 4); cudaMemset(a, 0, n * 4); cudaMemset(b, 0, n * 4);
  cudaFuncSetAttribute(chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaStream_t st; cudaStreamCreate(&st);
  cudaGraph_t g; cudaGraphExec_t ge;
  cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal);
  for (int i = 0; i < NL; ++i) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(threads); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    cudaLaunchKernelEx(&cfg, chain_kernel, (const float*)((i & 1) ? b : a), (i & 1) ? a : b, n, mode, early);
  }
  cudaStreamEndCapture(st, &g);
  cudaGraphInstantiate(&ge, g, 0);
  cudaGraphLaunch(ge, st); cudaStreamSynchronize(st);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int reps = 20;
  cudaEventRecord(e0, st);
  for (int r = 0; r < reps; ++r) cudaGraphLaunch(ge, st);
  cudaEventRecord(e1, st); cudaStreamSynchronize(st);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) printf("error: %s\n", cudaGetErrorString(e));
  cudaGraphExecDestroy(ge); cudaGraphDestroy(g); cudaStreamDestroy(st); cudaFree(a); cudaFree(b);
  return ms * 1e3f / (reps * NL);
}

int main() {
  const int NL = 128;
  const char* names[3] = {"empty", "wait+store", "wait+read 8KB+reduce+write"};
  printf("%-28s %5s %7s %7s %5s %4s  us/launch\n", "kernel", "grid", "threads", "smem", "early", "pdl");
  for (int mode = 0; mode < 3; ++mode)
    for (int cfg = 0; cfg < 4; ++cfg) {
      const int grid = cfg == 3 ? 296 : 148, threads = cfg >= 1 ? 320 : 576, smem = cfg == 0 ? 227 * 1024 : cfg == 1 ? 227 * 1024 : 100 * 1024;
      for (int pdl = 0; pdl < 2; ++pdl)
        for (int early = 0; early < 2; ++early) {
          if (!pdl && early) continue;
          printf("%-28s %5d %7d %7d %5d %4d  %.2f\n", names[mode], grid, threads, smem, early, pdl, run(grid, threads, smem, mode, early, pdl, NL));
        }
    }
  return 0;
}
