// Peer-memory plumbing for the tensor-parallel hand-off: a symmetric buffer per rank, allocated here with
// cudaMalloc (torch's caching allocator sub-allocates, and an IPC handle names a whole allocation), exported to the
// other ranks of the node as a CUDA IPC handle and mapped by them.  Setup-time only: nothing here runs per token.
// Replaces the NCCL all-reduce the reference issues from all_reduce_output (generate/tp.py:73-74) on the decode path;
// the data movement itself is in the GEMV kernel (csrc/qgemv_stream.cu: peer stores in the epilogue of the
// row-parallel Linear, arrival flags, the sum in the prologue of the next Linear).
#include "common.cuh"
#include <cstring>

using namespace b200;

namespace b200 {
// Consumer side of an ALL-GATHER over the hand-off slots (lm_head sharded by rows): slot r holds rank r's
// (batch, n_per_rank) outputs as self-validating packets `tag << 16 | bf16` written by that rank's GEMV epilogue; poll until the
// tag matches, unpack into out (batch, world * n_per_rank) bf16; the last CTA bumps the buffer's sequence counter.
__global__ void __launch_bounds__(256) tp_gather_kernel(const uint32_t* __restrict__ slots, int64_t slot_stride, int world, int batch,
                                                        int64_t npr, uint32_t* recv_seq, uint32_t* recv_done, uint32_t* err,
                                                        __nv_bfloat16* __restrict__ out) {
  griddep_launch();
  griddep_wait();
  const uint32_t tag = (__ldcg(recv_seq) % 65535u) + 1u;
  const int64_t chunks_per_row = npr >> 3, total = (int64_t)world * batch * chunks_per_row;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256) {
    const int r = (int)(i / (batch * chunks_per_row));
    const int64_t rem = i - (int64_t)r * batch * chunks_per_row;
    const int b = (int)(rem / chunks_per_row);
    const int64_t c = rem - (int64_t)b * chunks_per_row;
    const uint32_t* sp = slots + r * slot_stride + (int64_t)b * npr + c * 8;
    uint4 pa, pb;
    const long long t0 = clock64();
    for (;;) {
      asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(pa.x), "=r"(pa.y), "=r"(pa.z), "=r"(pa.w) : "l"(sp) : "memory");
      asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(pb.x), "=r"(pb.y), "=r"(pb.z), "=r"(pb.w) : "l"(sp + 4) : "memory");
      const bool ok = (pa.x >> 16) == tag && (pa.y >> 16) == tag && (pa.z >> 16) == tag && (pa.w >> 16) == tag &&
                      (pb.x >> 16) == tag && (pb.y >> 16) == tag && (pb.z >> 16) == tag && (pb.w >> 16) == tag;
      if (ok) break;
      if (clock64() - t0 > 4000000000LL) { if (err) *reinterpret_cast<volatile uint32_t*>(err) = 2u; break; }   // a peer died: error word, no hang
    }
    uint4 o;
    o.x = (pa.x & 0xffffu) | (pa.y << 16); o.y = (pa.z & 0xffffu) | (pa.w << 16);
    o.z = (pb.x & 0xffffu) | (pb.y << 16); o.w = (pb.z & 0xffffu) | (pb.w << 16);
    *reinterpret_cast<uint4*>(out + ((int64_t)b * world + r) * npr + c * 8) = o;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t old = atomicAdd(recv_done, 1u);
    if (old == gridDim.x - 1u) {
      *recv_done = 0u;
      *recv_seq = __ldcg(recv_seq) + 1u;
    }
  }
}
}  // namespace b200

extern "C" {

int b200_tp_gather(const b200_tp_handoff* h, int batch, int64_t n_per_rank, void* out, void* error_word, b200_stream_t stream) {
  B200_CHECK_ARG(h && out && h->slots && h->recv_seq && h->recv_done, "b200_tp_gather: null argument");
  B200_CHECK_ARG(h->world >= 2 && h->world <= 8 && batch >= 1 && batch <= 16 && n_per_rank > 0 && n_per_rank % 8 == 0,
                 "b200_tp_gather: world %d / batch %d / n_per_rank %lld out of range (n_per_rank must be a multiple of 8)", h->world, batch, (long long)n_per_rank);
  B200_CHECK_ARG(h->slot_stride >= (int64_t)batch * n_per_rank * 4 && h->slot_stride % 16 == 0 && (reinterpret_cast<uintptr_t>(h->slots) & 15) == 0 &&
                     (reinterpret_cast<uintptr_t>(out) & 15) == 0, "b200_tp_gather: slots too small or misaligned");
  const int64_t total = (int64_t)h->world * batch * (n_per_rank / 8);
  int64_t blocks = (total + 255) / 256;
  if (blocks > sm_count()) blocks = sm_count();
  B200_CUDA(launch(tp_gather_kernel, dim3((unsigned)blocks), dim3(256), 0, (cudaStream_t)stream, reinterpret_cast<const uint32_t*>(h->slots),
                   h->slot_stride / 4, h->world, batch, n_per_rank, reinterpret_cast<uint32_t*>(h->recv_seq), reinterpret_cast<uint32_t*>(h->recv_done),
                   reinterpret_cast<uint32_t*>(error_word), reinterpret_cast<__nv_bfloat16*>(out)));
  return B200_OK;
}

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
