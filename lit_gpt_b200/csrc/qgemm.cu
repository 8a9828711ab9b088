// Prefill GEMM entry point.  The tcgen05/TMEM implementation lives in qgemm_tcgen05.cu once enabled; until then
// this entry reports UNSUPPORTED so callers fail loudly instead of silently taking another path.
#include "common.cuh"

extern "C" {

size_t b200_qgemm_workspace_bytes(int64_t M, int64_t N, int64_t K) {
  (void)M; (void)N; (void)K;
  return 0;
}

int b200_qgemm(const b200_qgemm_args* a, b200_stream_t stream) {
  (void)stream;
  B200_CHECK_ARG(a != nullptr, "b200_qgemm: null args");
  b200::set_error("b200_qgemm: tcgen05 prefill GEMM not built into this library yet");
  return B200_ERR_UNSUPPORTED;
}

}  // extern "C"
