// Library-wide pieces of the C ABI: version and the thread-local last-error string.
#include "b2_common.cuh"

static thread_local char g_err[512] = "";

void b2_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* b2_last_error(void) { return g_err; }

#define B2_VERSION 100  // 0.1.0
extern "C" int b2_version(void) { return B2_VERSION; }

// 1 when a CUDA device is usable from this process, else 0 (never throws, never initialises torch).
extern "C" int b2_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}
