#include "host_common.h"

#include <cstring>

namespace b2d {

static thread_local char t_err[512] = "";
std::atomic<uint64_t> g_launch_count{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(t_err, sizeof(t_err), fmt, ap);
  va_end(ap);
}

void* driver_fn(const char* name) {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaError_t e = cudaGetDriverEntryPointByVersion(name, &fn, 12000, cudaEnableDefault, &q);
  if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || fn == nullptr) {
    set_error("driver entry point %s unavailable (%s)", name, cudaGetErrorString(e));
    (void)cudaGetLastError();
    return nullptr;
  }
  return fn;
}

}  // namespace b2d

extern "C" const char* b2d_last_error(void) { return b2d::t_err; }
extern "C" uint64_t b2d_launch_count(void) {
  return b2d::g_launch_count.load(std::memory_order_relaxed);
}
