// Library bookkeeping: error string, version, device query.
#include "common.cuh"
#include <stdarg.h>
#include <string.h>
#include <stdlib.h>

namespace fr {
static thread_local char g_err[1024] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
bool pdl_enabled() {
  static const bool on = [] { const char* e = getenv("FEDREC_PDL"); return !(e && atoi(e) == 0); }();
  return on;
}
int num_sms() {
  static int cache[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return kNumSMs;
  int n = __atomic_load_n(&cache[dev], __ATOMIC_RELAXED);
  if (n == 0) {
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = kNumSMs;
    __atomic_store_n(&cache[dev], n, __ATOMIC_RELAXED);
  }
  return n;
}
}  // namespace fr

FR_API const char* fr_last_error() { return fr::g_err; }
FR_API int fr_version() { return 100; }

// Fills {sm_count, cc_major, cc_minor, max_smem_optin} for the current device.
FR_API int fr_device_info(int* out4) {
  int dev = 0;
  FR_CHECK(cudaGetDevice(&dev));
  cudaDeviceProp p;
  FR_CHECK(cudaGetDeviceProperties(&p, dev));
  out4[0] = p.multiProcessorCount;
  out4[1] = p.major;
  out4[2] = p.minor;
  out4[3] = (int)p.sharedMemPerBlockOptin;
  return 0;
}
