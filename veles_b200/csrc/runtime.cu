// veles_b200 — tiny runtime shared by all kernel translation units:
// last-error string, device properties cache, version query.
#include "common.cuh"

#include <stdarg.h>
#include <string.h>

namespace vb {

static thread_local char g_last_error[1024] = "";

void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_last_error, sizeof(g_last_error), fmt, ap);
  va_end(ap);
}

int device_sm_count() {
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

}  // namespace vb

extern "C" const char* vb_last_error() { return vb::g_last_error; }

extern "C" int vb_version() { return 100; }

extern "C" int vb_sm_count() { return vb::device_sm_count(); }

// Compute-capability probe so the Python side can refuse to run the sm_100a
// kernels on anything else with a clear message instead of a launch failure.
extern "C" int vb_device_cc() {
  int dev = 0, major = 0, minor = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return -1;
  cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev);
  return major * 10 + minor;
}
