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

// SMs the persistent tcgen05 kernels may fill: all of them, minus the ones a pending data-parallel exchange kernel
// runs on (vb_set_sm_reserve).  Read at launch (= graph capture) time.
int g_pdl_enabled = 0;
static int g_sm_reserve = 0;
int gemm_sm_count() {
  int n = (device_sm_count() - g_sm_reserve) & ~1;   // whole TPCs: CTA pairs take two SMs
  return n < 2 ? 2 : n;
}

}  // namespace vb

// Keep `n` SMs free of persistent GEMM CTAs from now on (0 = none): the fused all-reduce + update kernel of the
// data-parallel group runs on them beside the backward pass.  Returns the previous value.
extern "C" int vb_set_sm_reserve(int n) {
  const int prev = vb::g_sm_reserve;
  vb::g_sm_reserve = n < 0 ? 0 : n;
  return prev;
}

// 1: kernels of this library are launched with programmatic stream serialisation from now on (each waits for its
// predecessor with griddepcontrol.wait, so results are unchanged); 0: ordinary launches.  Returns the previous value.
extern "C" int vb_set_pdl(int on) {
  const int prev = vb::g_pdl_enabled;
  vb::g_pdl_enabled = on ? 1 : 0;
  return prev;
}

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
