// How many thread-block clusters of a 576-thread, 227 KB CTA can be co-resident on this GPU?  (148 SMs, one CTA per SM:
// a grid of 148 CTAs in clusters of CS needs 148 / CS active clusters or it runs in two waves.)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/micro/cluster_occ tools/micro/cluster_occ.cu
#include <cuda_runtime.h>
#include <cstdio>
__global__ void __launch_bounds__(576, 1) k(float* p) { extern __shared__ float s[]; if (p) p[0] = s[0]; }
int main() {
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  cudaFuncSetAttribute(k, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  for (int cs : {1, 2, 4, 8, 16}) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(148 / cs * cs); cfg.blockDim = dim3(576); cfg.dynamicSmemBytes = 227 * 1024;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    int n = -1;
    cudaError_t e = cudaOccupancyMaxActiveClusters(&n, k, &cfg);
    printf("cluster size %2d: max active clusters %d (x%d = %d CTAs)  %s\n", cs, n, cs, n * cs, e == cudaSuccess ? "" : cudaGetErrorString(e));
  }
  return 0;
}
