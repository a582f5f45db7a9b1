// ABI bookkeeping: version, thread-local error string, device queries.
#include <cstring>

#include "common.cuh"

namespace b200attn {

static thread_local char g_err[512] = {0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int sm_count() {
  static thread_local int cached_dev = -1;
  static thread_local int cached = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
    cached = n;
    cached_dev = dev;
  }
  return cached;
}

}  // namespace b200attn

extern "C" {

int b200attn_abi_version(void) { return 21; }

const char* b200attn_last_error(void) { return b200attn::g_err; }

int b200attn_sm_count(void) { return b200attn::sm_count(); }

}  // extern "C"
