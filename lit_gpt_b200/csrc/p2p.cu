// Peer-memory plumbing for the tensor-parallel hand-off: a symmetric buffer per rank, allocated here with
// cudaMalloc (torch's caching allocator sub-allocates, and an IPC handle names a whole allocation), exported to the
// other ranks of the node as a CUDA IPC handle and mapped by them.  Setup-time only: nothing here runs per token.
// Replaces the NCCL all-reduce the reference issues from all_reduce_output (generate/tp.py:73-74) on the decode path;
// the data movement itself is in the GEMV kernel (csrc/qgemv_stream.cu: peer stores in the epilogue of the
// row-parallel Linear, arrival flags, the sum in the prologue of the next Linear).
#include "common.cuh"
#include <cstring>

using namespace b200;

extern "C" {

int b200_p2p_alloc(size_t bytes, void** ptr, void* handle64) {
  B200_CHECK_ARG(ptr && handle64 && bytes > 0, "b200_p2p_alloc: bad arguments");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  void* p = nullptr;
  B200_CUDA(cudaMalloc(&p, bytes));
  cudaError_t e = cudaMemset(p, 0, bytes);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  cudaIpcMemHandle_t h;
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    set_error("b200_p2p_alloc: %s", cudaGetErrorString(e));
    return B200_ERR_CUDA;
  }
  memcpy(handle64, &h, 64);
  *ptr = p;
  return B200_OK;
}

int b200_p2p_open(const void* handle64, void** ptr) {
  B200_CHECK_ARG(ptr && handle64, "b200_p2p_open: bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  void* p = nullptr;
  B200_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *ptr = p;
  return B200_OK;
}

int b200_p2p_close(void* ptr) {
  if (ptr) B200_CUDA(cudaIpcCloseMemHandle(ptr));
  return B200_OK;
}

int b200_p2p_free(void* ptr) {
  if (ptr) B200_CUDA(cudaFree(ptr));
  return B200_OK;
}

}  // extern "C"
