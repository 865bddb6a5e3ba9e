// Error plumbing and process-wide helpers of the b3d C ABI.
#include <stdarg.h>

#include "../../include/b3d.h"
#include "b3d_common.cuh"

static thread_local char g_err[1024] = "";

void b3d_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* b3d_last_error(void) { return g_err; }
extern "C" int b3d_version(void) { return 100; }

int b3d_num_sms() {
  static int sms[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (sms[dev] == 0) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
    sms[dev] = v;
  }
  return sms[dev];
}
