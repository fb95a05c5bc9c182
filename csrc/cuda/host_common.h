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

// Streaming multiprocessors of the current device (cudaDevAttrMultiProcessorCount, cached per device).
// Grids and tile schedules are sized from this at run time; on a GPU-less build host it reports the
// B200's 148 so that geometry queries still answer.
int sm_count();

// Programmatic dependent launch: every kernel of this library is launched with the
// programmatic-stream-serialization attribute and begins with griddepcontrol.wait /
// griddepcontrol.launch_dependents (pdl_wait / pdl_trigger in ptx.cuh), so kernel N+1 is
// scheduled and runs its prologue while kernel N is still executing; its wait returns once
// kernel N has completed and flushed.  B2D_PDL=0 turns the attribute off (the device-side
// instructions are then no-ops).  Captured into CUDA graphs the edges stay programmatic.
bool pdl_enabled();
// Every kernel asks for the same (maximal) shared-memory carve-out, so the SMs never have to
// re-partition L1/shared memory between consecutive kernels of a step (that switch drains the
// SM).  B2D_CARVEOUT=0 leaves the driver's per-kernel default.
void set_uniform_carveout(const void* kernel);

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                              cudaStream_t stream, Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid, cfg.blockDim = block, cfg.dynamicSmemBytes = smem, cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  set_uniform_carveout(reinterpret_cast<const void*>(kernel));
  count_launch();
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

}  // namespace b2d
