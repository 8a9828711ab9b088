// Library-level state of libb200gpt: error string, launch counter, PDL switch, device info.
#include "common.cuh"

namespace b200 {

static thread_local char g_err[512] = "";
std::atomic<int64_t> g_launches{0};
std::atomic<int> g_pdl{1};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

}  // namespace b200

extern "C" {

int b200_abi_version(void) { return B200GPT_ABI_VERSION; }

const char* b200_last_error(void) { return b200::g_err; }

int b200_device_info(int* sm_count, int* cc_major, int* cc_minor) {
  int dev = 0;
  B200_CUDA(cudaGetDevice(&dev));
  int n = 0, maj = 0, min = 0;
  B200_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
  B200_CUDA(cudaDeviceGetAttribute(&maj, cudaDevAttrComputeCapabilityMajor, dev));
  B200_CUDA(cudaDeviceGetAttribute(&min, cudaDevAttrComputeCapabilityMinor, dev));
  if (sm_count) *sm_count = n;
  if (cc_major) *cc_major = maj;
  if (cc_minor) *cc_minor = min;
  return B200_OK;
}

int b200_set_pdl(int enabled) { return b200::g_pdl.exchange(enabled ? 1 : 0); }

int64_t b200_launch_count(void) { return b200::g_launches.load(); }

}  // extern "C"
