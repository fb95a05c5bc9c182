#include "host_common.h"

#include <cstdlib>
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

bool pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("B2D_PDL");
    return e == nullptr || e[0] != '0';
  }();
  return on;
}

void set_uniform_carveout(const void* kernel) {
  static const bool on = [] {
    const char* e = getenv("B2D_CARVEOUT");
    return e == nullptr || e[0] != '0';
  }();
  if (!on) return;
  // small fixed table: the library has a few dozen kernels; lookups are lock-free reads
  static std::atomic<const void*> seen[256];
  for (auto& slot : seen) {
    const void* cur = slot.load(std::memory_order_acquire);
    if (cur == kernel) return;
    if (cur == nullptr) {
      const void* expected = nullptr;
      if (slot.compare_exchange_strong(expected, kernel)) {
        cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                             cudaSharedmemCarveoutMaxShared);
        (void)cudaGetLastError();
        return;
      }
      if (expected == kernel) return;
    }
  }
}

// Only used where no device can be queried (the GPU-less build host answering geometry queries).
constexpr int kFallbackSmCount = 148;   // B200

int sm_count() {
  static std::atomic<int> cached[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
    (void)cudaGetLastError();
    return kFallbackSmCount;
  }
  int n = cached[dev].load(std::memory_order_relaxed);
  if (n == 0) {
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) {
      (void)cudaGetLastError();
      n = kFallbackSmCount;
    }
    cached[dev].store(n, std::memory_order_relaxed);
  }
  return n;
}

void* driver_fn(const char* name) {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaError_t e = cudaGetDriverEntryPointByVersion(name, &fn, 12080, cudaEnableDefault, &q);
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
