// Error plumbing, version and device attribute cache of libfdsm.
#include <stdarg.h>

#include <mutex>

#include "fdsm_common.cuh"

namespace fdsm {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_status(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return FDSM_OK;
  set_error("%s: CUDA error %d (%s)", what, (int)e, cudaGetErrorString(e));
  return (int)e;
}

// Immutable per-process attribute cache (device of first use; one process per GPU).
int sm_count() {
  static int n = 0;
  static std::once_flag once;
  std::call_once(once, [] {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
  });
  return n;
}

}  // namespace fdsm

extern "C" int fdsm_version(void) { return FDSM_ABI_VERSION; }
extern "C" const char* fdsm_last_error_string(void) { return fdsm::g_err; }
extern "C" int fdsm_device_supported(void) {
  int dev = 0, major = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return 0;
  return major == 10 ? 1 : 0;
}
