// executor.cpp — native step executor: CUDA-graph capture of one training step and the
// host-side step loop that feeds it.
//
// The Python engine builds a step as a fixed list of kernel launches over fixed buffers.  The
// executor captures that list once per input staging slot (two slots: ping-pong) into CUDA
// graphs and then runs the hot loop natively:
//
//   copy stream : wait(slot consumed) -> memcpyAsync(x slice, y slice -> slot) -> record(ready)
//   main stream : wait(ready) -> cudaGraphLaunch(step graph of slot) -> record(consumed)
//                 -> memcpyAsync(metrics -> pinned ring)                     (device -> host)
//
// so the copy of step t+1 overlaps the compute of step t, there is no Python in the loop and
// no host synchronisation until the caller asks for it.  The source of a slice can be pinned
// host memory (end-to-end path) or a device-resident dataset (device-timed path); the copy
// direction is inferred.  This is the "fit" hot loop of README.md:304 / 392.
#include <cstring>
#include <new>

#include "../cuda/host_common.h"

using namespace b2d;

struct b2d_executor {
  cudaStream_t main = nullptr, copy = nullptr, side = nullptr;
  cudaEvent_t fork_ev = nullptr, join_ev = nullptr;
  cudaGraph_t graph[2] = {nullptr, nullptr};
  cudaGraphExec_t exec[2] = {nullptr, nullptr};
  cudaEvent_t ready[2] = {nullptr, nullptr}, consumed[2] = {nullptr, nullptr};
  cudaEvent_t t0 = nullptr, t1 = nullptr;
  void* x_dev[2] = {nullptr, nullptr};
  void* y_dev[2] = {nullptr, nullptr};
  size_t x_bytes = 0, y_bytes = 0;
  const void* metrics_dev = nullptr;
  const void* metrics_dev_slot[2] = {nullptr, nullptr};  // weight ping-pong: where slot s leaves the sums
  void* metrics_ring = nullptr;  // pinned host, ring_len entries of metrics_bytes
  size_t metrics_bytes = 0;
  int ring_len = 0;
  uint64_t steps_done = 0;
  uint64_t nodes[2] = {0, 0};  // kernel nodes per captured graph
  bool capturing = false;
};

extern "C" int b2d_exec_create(b2d_executor** out) {
  B2D_REQUIRE(out != nullptr, "exec_create: null out");
  b2d_executor* e = new (std::nothrow) b2d_executor();
  B2D_REQUIRE(e != nullptr, "exec_create: out of host memory");
  B2D_CUDA_CHECK(cudaStreamCreateWithFlags(&e->main, cudaStreamNonBlocking));
  B2D_CUDA_CHECK(cudaStreamCreateWithFlags(&e->copy, cudaStreamNonBlocking));
  B2D_CUDA_CHECK(cudaStreamCreateWithFlags(&e->side, cudaStreamNonBlocking));
  B2D_CUDA_CHECK(cudaEventCreateWithFlags(&e->fork_ev, cudaEventDisableTiming));
  B2D_CUDA_CHECK(cudaEventCreateWithFlags(&e->join_ev, cudaEventDisableTiming));
  for (int i = 0; i < 2; ++i) {
    B2D_CUDA_CHECK(cudaEventCreateWithFlags(&e->ready[i], cudaEventDisableTiming));
    B2D_CUDA_CHECK(cudaEventCreateWithFlags(&e->consumed[i], cudaEventDisableTiming));
  }
  B2D_CUDA_CHECK(cudaEventCreate(&e->t0));
  B2D_CUDA_CHECK(cudaEventCreate(&e->t1));
  *out = e;
  return B2D_OK;
}

extern "C" void b2d_exec_destroy(b2d_executor* e) {
  if (e == nullptr) return;
  for (int i = 0; i < 2; ++i) {
    if (e->exec[i]) cudaGraphExecDestroy(e->exec[i]);
    if (e->graph[i]) cudaGraphDestroy(e->graph[i]);
    if (e->ready[i]) cudaEventDestroy(e->ready[i]);
    if (e->consumed[i]) cudaEventDestroy(e->consumed[i]);
  }
  if (e->t0) cudaEventDestroy(e->t0);
  if (e->t1) cudaEventDestroy(e->t1);
  if (e->main) cudaStreamDestroy(e->main);
  if (e->copy) cudaStreamDestroy(e->copy);
  if (e->side) cudaStreamDestroy(e->side);
  if (e->fork_ev) cudaEventDestroy(e->fork_ev);
  if (e->join_ev) cudaEventDestroy(e->join_ev);
  delete e;
}

extern "C" void* b2d_exec_stream(b2d_executor* e) { return e ? e->main : nullptr; }

// Side branch inside a step (bucketed communication overlapped with compute): work launched on
// b2d_exec_side_stream() between fork and join runs concurrently with the main stream; inside a
// capture the two calls become the fork / join edges of the step graph.
extern "C" void* b2d_exec_side_stream(b2d_executor* e) { return e ? e->side : nullptr; }
extern "C" int b2d_exec_fork(b2d_executor* e) {
  B2D_REQUIRE(e != nullptr, "exec_fork: null executor");
  B2D_CUDA_CHECK(cudaEventRecord(e->fork_ev, e->main));
  B2D_CUDA_CHECK(cudaStreamWaitEvent(e->side, e->fork_ev, 0));
  return B2D_OK;
}
extern "C" int b2d_exec_join(b2d_executor* e) {
  B2D_REQUIRE(e != nullptr, "exec_join: null executor");
  B2D_CUDA_CHECK(cudaEventRecord(e->join_ev, e->side));
  B2D_CUDA_CHECK(cudaStreamWaitEvent(e->main, e->join_ev, 0));
  return B2D_OK;
}

extern "C" int b2d_exec_set_io(b2d_executor* e, int slot, void* x_dev, size_t x_bytes, void* y_dev,
                               size_t y_bytes) {
  B2D_REQUIRE(e != nullptr && (slot == 0 || slot == 1), "exec_set_io: bad slot");
  e->x_dev[slot] = x_dev, e->y_dev[slot] = y_dev;
  e->x_bytes = x_bytes, e->y_bytes = y_bytes;
  return B2D_OK;
}

extern "C" int b2d_exec_set_metrics(b2d_executor* e, const void* metrics_dev, size_t bytes,
                                    void* pinned_ring, int ring_len) {
  B2D_REQUIRE(e != nullptr && ring_len > 0, "exec_set_metrics: bad arguments");
  e->metrics_dev = metrics_dev, e->metrics_bytes = bytes;
  e->metrics_ring = pinned_ring, e->ring_len = ring_len;
  return B2D_OK;
}

// With weight ping-pong the running sums alternate between the two weight buffers: the step replayed
// from `slot` leaves them at `metrics_dev`.
extern "C" int b2d_exec_set_metrics_slot(b2d_executor* e, int slot, const void* metrics_dev) {
  B2D_REQUIRE(e != nullptr && (slot == 0 || slot == 1), "exec_set_metrics_slot: bad arguments");
  e->metrics_dev_slot[slot] = metrics_dev;
  return B2D_OK;
}

// Everything launched on b2d_exec_stream() between begin and end becomes the step graph.
extern "C" int b2d_exec_begin_capture(b2d_executor* e) {
  B2D_REQUIRE(e != nullptr && !e->capturing, "exec_begin_capture: bad state");
  B2D_CUDA_CHECK(cudaStreamBeginCapture(e->main, cudaStreamCaptureModeThreadLocal));
  e->capturing = true;
  return B2D_OK;
}

extern "C" int b2d_exec_end_capture(b2d_executor* e, int slot) {
  B2D_REQUIRE(e != nullptr && e->capturing && (slot == 0 || slot == 1), "exec_end_capture: bad state");
  e->capturing = false;
  cudaGraph_t g = nullptr;
  B2D_CUDA_CHECK(cudaStreamEndCapture(e->main, &g));
  if (e->exec[slot]) cudaGraphExecDestroy(e->exec[slot]);
  if (e->graph[slot]) cudaGraphDestroy(e->graph[slot]);
  e->graph[slot] = g;
  size_t n = 0;
  B2D_CUDA_CHECK(cudaGraphGetNodes(g, nullptr, &n));
  e->nodes[slot] = n;
  B2D_CUDA_CHECK(cudaGraphInstantiate(&e->exec[slot], g, 0));
  B2D_CUDA_CHECK(cudaGraphUpload(e->exec[slot], e->main));
  return B2D_OK;
}

extern "C" uint64_t b2d_exec_graph_nodes(const b2d_executor* e, int slot) {
  return e ? e->nodes[slot & 1] : 0;
}

// Runs n_steps steps.  Step t reads its inputs at src + ((first_step + t) % steps_in_src) *
// stride.  If elapsed_ms is non-null the loop is bracketed by CUDA events on the main stream
// and the call synchronises; otherwise it returns as soon as the work is enqueued unless
// `sync` is set.
extern "C" int b2d_exec_run(b2d_executor* e, const void* x_src, size_t x_stride, const void* y_src,
                            size_t y_stride, int64_t first_step, int64_t steps_in_src,
                            int64_t n_steps, int sync, float* elapsed_ms) {
  B2D_REQUIRE(e != nullptr && e->exec[0] != nullptr && e->exec[1] != nullptr,
              "exec_run: step graphs have not been captured");
  B2D_REQUIRE(steps_in_src > 0 && n_steps >= 0, "exec_run: bad step counts");
  if (elapsed_ms != nullptr) B2D_CUDA_CHECK(cudaEventRecord(e->t0, e->main));
  for (int64_t t = 0; t < n_steps; ++t) {
    const int slot = static_cast<int>((e->steps_done + t) & 1);
    const int64_t s = (first_step + t) % steps_in_src;
    // The slot may be overwritten once the step that last used it has finished.
    B2D_CUDA_CHECK(cudaStreamWaitEvent(e->copy, e->consumed[slot], 0));
    B2D_CUDA_CHECK(cudaMemcpyAsync(e->x_dev[slot], static_cast<const char*>(x_src) + s * x_stride,
                                   e->x_bytes, cudaMemcpyDefault, e->copy));
    B2D_CUDA_CHECK(cudaMemcpyAsync(e->y_dev[slot], static_cast<const char*>(y_src) + s * y_stride,
                                   e->y_bytes, cudaMemcpyDefault, e->copy));
    B2D_CUDA_CHECK(cudaEventRecord(e->ready[slot], e->copy));
    B2D_CUDA_CHECK(cudaStreamWaitEvent(e->main, e->ready[slot], 0));
    B2D_CUDA_CHECK(cudaGraphLaunch(e->exec[slot], e->main));
    B2D_CUDA_CHECK(cudaEventRecord(e->consumed[slot], e->main));
    if (e->metrics_ring != nullptr) {
      char* dst = static_cast<char*>(e->metrics_ring) +
                  ((e->steps_done + t) % e->ring_len) * e->metrics_bytes;
      const void* src = e->metrics_dev_slot[slot] != nullptr ? e->metrics_dev_slot[slot] : e->metrics_dev;
      B2D_CUDA_CHECK(cudaMemcpyAsync(dst, src, e->metrics_bytes, cudaMemcpyDeviceToHost, e->main));
    }
    count_launch(e->nodes[slot]);
  }
  e->steps_done += n_steps;
  if (elapsed_ms != nullptr) {
    B2D_CUDA_CHECK(cudaEventRecord(e->t1, e->main));
    B2D_CUDA_CHECK(cudaEventSynchronize(e->t1));
    B2D_CUDA_CHECK(cudaEventElapsedTime(elapsed_ms, e->t0, e->t1));
  } else if (sync) {
    B2D_CUDA_CHECK(cudaStreamSynchronize(e->main));
  }
  return B2D_OK;
}

extern "C" int b2d_exec_sync(b2d_executor* e) {
  B2D_REQUIRE(e != nullptr, "exec_sync: null executor");
  B2D_CUDA_CHECK(cudaStreamSynchronize(e->copy));
  B2D_CUDA_CHECK(cudaStreamSynchronize(e->main));
  return B2D_OK;
}

extern "C" uint64_t b2d_exec_steps_done(const b2d_executor* e) { return e ? e->steps_done : 0; }

// Pinned host allocations for the input pipeline and the metrics ring.
extern "C" int b2d_host_alloc(void** out, size_t bytes) {
  B2D_CUDA_CHECK(cudaHostAlloc(out, bytes, cudaHostAllocDefault));
  return B2D_OK;
}
extern "C" int b2d_host_free(void* p) {
  B2D_CUDA_CHECK(cudaFreeHost(p));
  return B2D_OK;
}
