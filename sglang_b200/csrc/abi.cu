// Error reporting and device queries of the C ABI (include/b200attn.h).
#include <cstring>

#include "common.cuh"

namespace b200 {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int sm_count() {
  static int cached = 0;
  if (cached == 0) {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) == cudaSuccess &&
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess)
      cached = n;
  }
  return cached;
}

}  // namespace b200

extern "C" const char* b200attn_last_error(void) { return b200::g_err; }

extern "C" int b200attn_abi_version(void) { return B200ATTN_ABI_VERSION; }

extern "C" int b200attn_device_info(int* sm_count, int* cc_major, int* cc_minor) {
  using namespace b200;
  int dev = 0, n = 0, major = 0, minor = 0;
  B200_CHECK_CUDA(cudaGetDevice(&dev));
  B200_CHECK_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
  B200_CHECK_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  B200_CHECK_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev));
  if (sm_count) *sm_count = n;
  if (cc_major) *cc_major = major;
  if (cc_minor) *cc_minor = minor;
  if (major != 10) {
    set_error("device is sm_%d%d; libb200attn is built for sm_100a only", major, minor);
    return B200_ENOTSM100;
  }
  return B200_OK;
}
