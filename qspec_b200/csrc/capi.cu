// Library-wide C-ABI plumbing: status text, launch counter, device queries.
#include "common.cuh"
#include <stdlib.h>

namespace qspec {
static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

bool pdl_enabled() {
  static int cached = -1;
  if (cached < 0) {
    const char* v = getenv("QSPEC_PDL");
    cached = (v && v[0] == '0') ? 0 : 1;
  }
  return cached != 0;
}

int sm_count() {
  static int cached = -1;
  if (cached < 0) {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) == cudaSuccess &&
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess)
      cached = n;
    else
      return 0;
  }
  return cached;
}
}  // namespace qspec

extern "C" int qspec_abi_version(void) { return 1; }
extern "C" const char* qspec_last_error(void) { return qspec::g_err; }
extern "C" uint64_t qspec_launch_count(void) { return qspec::g_launches.load(); }
extern "C" int qspec_sm_count(void) { return qspec::sm_count(); }
