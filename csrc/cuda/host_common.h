// host_common.h — error reporting, launch accounting and driver entry points shared by the
// host side of every .cu file in libb200dist_cuda.so.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "../../include/b200dist.h"

namespace b2d {

// Thread-local message returned by b2d_last_error().
void set_error(const char* fmt, ...) __attribute__((format(printf, 1, 2)));
extern std::atomic<uint64_t> g_launch_count;
inline void count_launch(uint64_t n = 1) { g_launch_count.fetch_add(n, std::memory_order_relaxed); }

// Driver API entry points are resolved at run time through the runtime (no link-time
// dependency on libcuda, so the library loads on GPU-less build hosts).
void* driver_fn(const char* name);

#define B2D_CUDA_CHECK(expr)                                                              \
  do {                                                                                    \
    cudaError_t _e = (expr);                                                              \
    if (_e != cudaSuccess) {                                                              \
      ::b2d::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, \
                       __LINE__);                                                         \
      return B2D_ERR_CUDA;                                                                \
    }                                                                                     \
  } while (0)

#define B2D_CU_CHECK(expr)                                                                 \
  do {                                                                                     \
    CUresult _r = (expr);                                                                  \
    if (_r != CUDA_SUCCESS) {                                                              \
      ::b2d::set_error("%s failed: CUresult %d (%s:%d)", #expr, (int)_r, __FILE__, __LINE__); \
      return B2D_ERR_CUDA;                                                                 \
    }                                                                                      \
  } while (0)

#define B2D_REQUIRE(cond, ...)        \
  do {                                \
    if (!(cond)) {                    \
      ::b2d::set_error(__VA_ARGS__);  \
      return B2D_ERR_INVALID;         \
    }                                 \
  } while (0)

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

}  // namespace b2d
