// Error plumbing and device queries shared by all entry points.
#include "common.cuh"

#include <cstdlib>

namespace nfb {

static thread_local char g_error[512];

char* error_buffer() { return g_error; }

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
  return code;
}

int launch_status(const char* what) {
  cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) {
    cudaGetLastError();
    snprintf(g_error, sizeof(g_error), "%s: %s", what, cudaGetErrorString(e));
    return static_cast<int>(e);
  }
  return 0;
}

bool pdl_enabled() {
  static const bool on = [] {
    const char* v = getenv("NFB_PDL");   // opt-in: measured slower than plain graph edges
    return v && v[0] == '1';
  }();
  return on;
}

int sm_count() {
  static thread_local int cached_dev = -1, cached = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    cudaDeviceGetAttribute(&cached, cudaDevAttrMultiProcessorCount, dev);
    cached_dev = dev;
    if (cached <= 0) cached = 148;
  }
  return cached;
}

}  // namespace nfb

extern "C" int nfb_version(void) { return 100; }
extern "C" const char* nfb_last_error(void) { return nfb::error_buffer(); }
