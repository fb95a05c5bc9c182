// ops.cpp — C++ CPU implementations of every op in the training step (libb200dist_cpu.so).
//
// They serve three purposes: the README-faithful CPU run
// (the reference itself ran on CPU workers, README.md:398), the timed CPU baseline, and the
// fp32 numerical oracle for the sm_100a kernels.  Layouts match the CUDA side exactly:
// NHWC activations, HWIO conv kernels, (in, out) Dense kernels.  Loops are written so the
// innermost dimension is contiguous (the compiler vectorises them); work is split across a
// small std::thread pool by output rows.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <memory>
#include <pthread.h>
#include <sched.h>
#include <unistd.h>
#include <cstdlib>
#include <condition_variable>
#include <mutex>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>
#include <thread>
#include <vector>

#define B2C_API extern "C" __attribute__((visibility("default")))

namespace {

int g_threads = 0;
int g_cpu_offset = -1;   // index (into the affinity mask) of the CPU the calling thread's share starts at

int num_threads() {
  if (g_threads > 0) return g_threads;
  unsigned hw = std::thread::hardware_concurrency();
  return static_cast<int>(std::max(1u, std::min(hw, 8u)));
}

// Persistent worker pool: threads are created once and parked on a condition variable; a parallel_for posts
// one job (a range cut into chunks handed out by an atomic counter), the caller works too, and returns when
// every chunk is done.  (The first version spawned and joined std::threads per call: ~0.1 ms per op.)
class Pool {
 public:
  static Pool& get() {
    static Pool p;
    return p;
  }
  // A job is finished when every CHUNK is done, not when every worker has checked in: a worker that wakes
  // late finds no chunk left and goes back to sleep, so a short job never waits for the slowest wake-up.
  struct Job {
    const std::function<void(int64_t, int64_t)>* fn;
    int64_t n, chunk, chunks;
    std::atomic<int64_t> next{0}, done{0};
  };
  void run(int64_t n, int64_t chunk, int want, const std::function<void(int64_t, int64_t)>& fn) {
    std::lock_guard<std::mutex> serial(run_mu_);          // one job at a time
    ensure(want - 1);
    auto job = std::make_shared<Job>();
    job->fn = &fn, job->n = n, job->chunk = chunk, job->chunks = (n + chunk - 1) / chunk;
    {
      std::lock_guard<std::mutex> lk(mu_);
      job_ = job;
      active_ = std::min<int>(want - 1, (int)workers_.size());
    }
    generation_.fetch_add(1, std::memory_order_release);
    cv_.notify_all();
    work(*job);
    for (int spin = 0; job->done.load(std::memory_order_acquire) < job->chunks; ++spin) {
      if (spin < 4096) __builtin_ia32_pause();
      else std::this_thread::yield();
    }
  }

 private:
  Pool() = default;
  ~Pool() {
    {
      std::lock_guard<std::mutex> lk(mu_);
      stop_ = true;
    }
    generation_.fetch_add(1, std::memory_order_release);
    cv_.notify_all();
    for (auto& t : workers_) t.join();
  }
  void ensure(int k) {
    while ((int)workers_.size() < k) {
      const int id = (int)workers_.size();
      workers_.emplace_back([this, id] { loop(id); });
      pin(workers_.back(), id);
    }
  }
  // Workers are woken by one thread and, left to the scheduler's wake-affine placement, all land on the
  // waker's CPU for bursts as short as these (measured: 8 threads on 1 CPU).  Each worker is therefore pinned
  // to its own CPU of the process's affinity mask, starting at a per-process offset so that several worker
  // processes on one host do not stack on the same cores.  B200DIST_CPU_PIN=0 leaves placement to the OS.
  static void pin(std::thread& t, int id) {
    const char* e = std::getenv("B200DIST_CPU_PIN");
    if (e != nullptr && e[0] == '0') return;
    cpu_set_t allowed;
    CPU_ZERO(&allowed);
    if (sched_getaffinity(0, sizeof(allowed), &allowed) != 0) return;
    std::vector<int> cpus;
    for (int c = 0; c < CPU_SETSIZE; ++c)
      if (CPU_ISSET(c, &allowed)) cpus.push_back(c);
    if (cpus.size() < 2) return;
    const size_t off = g_cpu_offset >= 0 ? static_cast<size_t>(g_cpu_offset) % cpus.size()
                                         : static_cast<size_t>(getpid()) % cpus.size();
    cpu_set_t one;
    CPU_ZERO(&one);
    CPU_SET(cpus[(off + id + 1) % cpus.size()], &one);
    pthread_setaffinity_np(t.native_handle(), sizeof(one), &one);
  }
  static void work(Job& j) {
    for (;;) {
      const int64_t b = j.next.fetch_add(j.chunk, std::memory_order_relaxed);
      if (b >= j.n) break;
      (*j.fn)(b, std::min(j.n, b + j.chunk));      // the caller is still inside run(): fn is alive
      j.done.fetch_add(1, std::memory_order_release);
    }
  }
  void loop(int id) {
    // After a job a worker keeps polling for the next one for a short while (the ops of a training step
    // follow each other within tens of microseconds) before it parks on the condition variable.
    static const int spin_us = [] {
      const char* e = std::getenv("B200DIST_CPU_SPIN_US");
      return e != nullptr ? std::max(0, atoi(e)) : 200;
    }();
    uint64_t seen = 0;
    for (;;) {
      const auto t0 = std::chrono::steady_clock::now();
      while (generation_.load(std::memory_order_acquire) == seen) {
        if (std::chrono::steady_clock::now() - t0 > std::chrono::microseconds(spin_us)) {
          std::unique_lock<std::mutex> lk(mu_);
          cv_.wait(lk, [&] { return generation_.load(std::memory_order_acquire) != seen; });
          break;
        }
        __builtin_ia32_pause();
      }
      std::shared_ptr<Job> job;
      {
        std::lock_guard<std::mutex> lk(mu_);
        if (stop_) return;
        seen = generation_.load(std::memory_order_acquire);
        if (id < active_) job = job_;
      }
      if (job) work(*job);
    }
  }
  std::mutex mu_, run_mu_;
  std::condition_variable cv_;
  std::vector<std::thread> workers_;
  std::shared_ptr<Job> job_;
  std::atomic<uint64_t> generation_{0};
  int active_ = 0;
  bool stop_ = false;
};

// Parallel for over [0, n) in chunks of >= grain; runs inline when the range is small.
void parallel_for(int64_t n, int64_t grain, const std::function<void(int64_t, int64_t)>& fn) {
  if (n <= 0) return;
  const int nt = static_cast<int>(std::min<int64_t>(num_threads(), (n + grain - 1) / grain));
  if (nt <= 1) {
    fn(0, n);
    return;
  }
  // ~4 chunks per thread for balance, never smaller than the grain
  const int64_t chunk = std::max<int64_t>(grain, (n + 4 * nt - 1) / (4 * nt));
  Pool::get().run(n, chunk, nt, fn);
}

struct ConvShape {
  int N, H, W, Ci, Co, KH, KW, OH, OW, sh, sw, ph, pw;
};

// ph, pw = padding at the top / left; OH, OW = output size (<= 0: derive from symmetric padding).
// TensorFlow's 'same' padding is asymmetric for even totals, which the explicit size expresses.
ConvShape make_shape(int N, int H, int W, int Ci, int Co, int KH, int KW, int sh, int sw, int ph,
                     int pw, int OH, int OW) {
  return ConvShape{N, H, W, Ci, Co, KH, KW, OH > 0 ? OH : (H + 2 * ph - KH) / sh + 1,
                   OW > 0 ? OW : (W + 2 * pw - KW) / sw + 1, sh, sw, ph, pw};
}

}  // namespace

B2C_API void b2c_set_num_threads(int n) { g_threads = n; }
// Several worker processes on one host: give each `threads` CPUs starting at index `first_cpu` of the
// affinity mask (call before the first op; the worker pool pins its threads when it creates them).
B2C_API void b2c_set_thread_layout(int threads, int first_cpu) { g_threads = threads, g_cpu_offset = first_cpu; }

B2C_API void b2c_gemm(const float* A, int64_t sam, int64_t sak, const float* B, int64_t sbk, int64_t sbn,
                      float* D, int64_t ldd, const float* bias, int M, int N, int K, int relu, int accumulate);

namespace {
// Convolutions with a reduction of at least kGemmConvK go through a patch matrix and the tiled GEMM
// (col[pixel][(kh, kw, ci)] matches the row order of the HWIO filter viewed as [KH*KW*Ci][Co]); the tiny
// first-layer filters (README: K = 9) keep the direct loops below.
constexpr int kGemmConvK = 32;

void im2col(const float* x, float* col, const ConvShape& s) {
  const int K = s.KH * s.KW * s.Ci;
  parallel_for((int64_t)s.N * s.OH, 2, [&](int64_t b, int64_t e) {
    for (int64_t r = b; r < e; ++r) {
      const int n = r / s.OH, oh = r % s.OH;
      for (int ow = 0; ow < s.OW; ++ow) {
        float* dst = col + ((r * s.OW) + ow) * (int64_t)K;
        for (int kh = 0; kh < s.KH; ++kh) {
          const int ih = oh * s.sh - s.ph + kh;
          for (int kw = 0; kw < s.KW; ++kw) {
            const int iw = ow * s.sw - s.pw + kw;
            float* d = dst + (int64_t)(kh * s.KW + kw) * s.Ci;
            if (ih < 0 || ih >= s.H || iw < 0 || iw >= s.W) {
              for (int ci = 0; ci < s.Ci; ++ci) d[ci] = 0.f;
            } else {
              const float* src = x + (((int64_t)n * s.H + ih) * s.W + iw) * s.Ci;
              for (int ci = 0; ci < s.Ci; ++ci) d[ci] = src[ci];
            }
          }
        }
      }
    }
  });
}

// dx[n, ih, iw, :] += dcol entries of every (output pixel, tap) that read it; one image per task.
void col2im(const float* dcol, float* dx, const ConvShape& s) {
  const int K = s.KH * s.KW * s.Ci;
  parallel_for(s.N, 1, [&](int64_t b, int64_t e) {
    for (int64_t n = b; n < e; ++n) {
      float* img = dx + n * (int64_t)s.H * s.W * s.Ci;
      for (int64_t i = 0; i < (int64_t)s.H * s.W * s.Ci; ++i) img[i] = 0.f;
      for (int oh = 0; oh < s.OH; ++oh)
        for (int ow = 0; ow < s.OW; ++ow) {
          const float* src = dcol + (((n * s.OH) + oh) * s.OW + ow) * (int64_t)K;
          for (int kh = 0; kh < s.KH; ++kh) {
            const int ih = oh * s.sh - s.ph + kh;
            if (ih < 0 || ih >= s.H) continue;
            for (int kw = 0; kw < s.KW; ++kw) {
              const int iw = ow * s.sw - s.pw + kw;
              if (iw < 0 || iw >= s.W) continue;
              float* d = img + ((int64_t)ih * s.W + iw) * s.Ci;
              const float* g = src + (int64_t)(kh * s.KW + kw) * s.Ci;
              for (int ci = 0; ci < s.Ci; ++ci) d[ci] += g[ci];
            }
          }
        }
    }
  });
}
}  // namespace

// y[n,oh,ow,:] = act(b + sum_{kh,kw,ci} x[n,ih,iw,ci] * w[kh,kw,ci,:])               [K1]
B2C_API void b2c_conv2d_fwd(const float* x, const float* w, const float* bias, float* y, int N, int H,
                            int W, int Ci, int Co, int KH, int KW, int sh, int sw, int ph, int pw,
                            int OH, int OW, int relu) {
  const ConvShape s = make_shape(N, H, W, Ci, Co, KH, KW, sh, sw, ph, pw, OH, OW);
  if (KH * KW * Ci >= kGemmConvK) {
    const int K = KH * KW * Ci;
    const int64_t pix = (int64_t)N * s.OH * s.OW;
    std::vector<float> col((size_t)pix * K);
    im2col(x, col.data(), s);
    b2c_gemm(col.data(), K, 1, w, Co, 1, y, Co, bias, (int)pix, Co, K, relu, 0);
    return;
  }
  parallel_for((int64_t)N * s.OH, 4, [&](int64_t b, int64_t e) {
    std::vector<float> acc(Co);
    for (int64_t r = b; r < e; ++r) {
      const int n = r / s.OH, oh = r % s.OH;
      for (int ow = 0; ow < s.OW; ++ow) {
        for (int c = 0; c < Co; ++c) acc[c] = bias ? bias[c] : 0.f;
        for (int kh = 0; kh < KH; ++kh) {
          const int ih = oh * sh - ph + kh;
          if (ih < 0 || ih >= H) continue;
          for (int kw = 0; kw < KW; ++kw) {
            const int iw = ow * sw - pw + kw;
            if (iw < 0 || iw >= W) continue;
            const float* xp = x + (((int64_t)n * H + ih) * W + iw) * Ci;
            const float* wp = w + (int64_t)(kh * KW + kw) * Ci * Co;
            for (int ci = 0; ci < Ci; ++ci) {
              const float xv = xp[ci];
              const float* wr = wp + (int64_t)ci * Co;
              for (int c = 0; c < Co; ++c) acc[c] += xv * wr[c];
            }
          }
        }
        float* yp = y + (((int64_t)n * s.OH + oh) * s.OW + ow) * Co;
        for (int c = 0; c < Co; ++c) yp[c] = relu ? std::max(acc[c], 0.f) : acc[c];
      }
    }
  });
}

// dW[kh,kw,ci,:] += sum x[n,ih,iw,ci] * dy[n,oh,ow,:] ;  db[:] += sum dy            [K9]
B2C_API void b2c_conv2d_wgrad(const float* x, const float* dy, float* dw, float* db, int N, int H,
                              int W, int Ci, int Co, int KH, int KW, int sh, int sw, int ph, int pw,
                              int OH, int OW) {
  const ConvShape s = make_shape(N, H, W, Ci, Co, KH, KW, sh, sw, ph, pw, OH, OW);
  const int64_t taps = (int64_t)KH * KW * Ci;
  if (taps >= kGemmConvK) {
    // dW[K, Co] += col^T . dy ; db += column sums of dy
    const int K = (int)taps;
    const int64_t pix = (int64_t)N * s.OH * s.OW;
    std::vector<float> col((size_t)pix * K);
    im2col(x, col.data(), s);
    b2c_gemm(col.data(), 1, K, dy, Co, 1, dw, Co, nullptr, K, Co, (int)pix, 0, 1);
    if (db != nullptr)
      for (int64_t r = 0; r < pix; ++r)
        for (int c = 0; c < Co; ++c) db[c] += dy[r * Co + c];
    return;
  }
  if (taps * Co <= 16384 && N >= 2) {
    // Small filter (e.g. the README's 3x3x1x32): the reduction runs over many pixels and few taps, so the
    // work is split over fixed groups of images, each into its own partial filter, and the partials are
    // added in group order afterwards — the result does not depend on the number of threads.
    constexpr int kGroup = 4;
    const int groups = (N + kGroup - 1) / kGroup;
    std::vector<float> part((size_t)groups * (taps * Co + Co), 0.f);
    parallel_for(groups, 1, [&](int64_t g0, int64_t g1) {
      for (int64_t g = g0; g < g1; ++g) {
        float* __restrict__ pw_ = part.data() + g * (taps * Co + Co);
        float* __restrict__ pb = pw_ + taps * Co;
        for (int n = (int)g * kGroup; n < std::min(N, (int)(g + 1) * kGroup); ++n)
          for (int oh = 0; oh < s.OH; ++oh)
            for (int ow = 0; ow < s.OW; ++ow) {
              const float* __restrict__ dp = dy + (((int64_t)n * s.OH + oh) * s.OW + ow) * Co;
              if (db != nullptr)
                for (int c = 0; c < Co; ++c) pb[c] += dp[c];
              for (int kh = 0; kh < KH; ++kh) {
                const int ih = oh * sh - ph + kh;
                if (ih < 0 || ih >= H) continue;
                for (int kw = 0; kw < KW; ++kw) {
                  const int iw = ow * sw - pw + kw;
                  if (iw < 0 || iw >= W) continue;
                  const float* __restrict__ xp = x + (((int64_t)n * H + ih) * W + iw) * Ci;
                  float* __restrict__ dwr = pw_ + (int64_t)(kh * KW + kw) * Ci * Co;
                  for (int ci = 0; ci < Ci; ++ci) {
                    const float xv = xp[ci];
                    for (int c = 0; c < Co; ++c) dwr[ci * Co + c] += xv * dp[c];
                  }
                }
              }
            }
      }
    });
    for (int g = 0; g < groups; ++g) {
      const float* pg = part.data() + (int64_t)g * (taps * Co + Co);
      for (int64_t i = 0; i < taps * Co; ++i) dw[i] += pg[i];
      if (db != nullptr)
        for (int c = 0; c < Co; ++c) db[c] += pg[taps * Co + c];
    }
    return;
  }
  // Parallel over filter taps (kh, kw, ci): each owns a disjoint dW row, so no reduction.
  parallel_for(taps, 1, [&](int64_t b, int64_t e) {
    for (int64_t tap = b; tap < e; ++tap) {
      const int ci = tap % Ci, kw = (tap / Ci) % KW, kh = tap / ((int64_t)Ci * KW);
      float* dwr = dw + tap * Co;
      for (int n = 0; n < N; ++n)
        for (int oh = 0; oh < s.OH; ++oh) {
          const int ih = oh * sh - ph + kh;
          if (ih < 0 || ih >= H) continue;
          for (int ow = 0; ow < s.OW; ++ow) {
            const int iw = ow * sw - pw + kw;
            if (iw < 0 || iw >= W) continue;
            const float xv = x[(((int64_t)n * H + ih) * W + iw) * Ci + ci];
            const float* dp = dy + (((int64_t)n * s.OH + oh) * s.OW + ow) * Co;
            for (int c = 0; c < Co; ++c) dwr[c] += xv * dp[c];
          }
        }
    }
  });
  if (db != nullptr) {
    const int64_t rows = (int64_t)N * s.OH * s.OW;
    for (int64_t r = 0; r < rows; ++r)
      for (int c = 0; c < Co; ++c) db[c] += dy[r * Co + c];
  }
}

// dx[n,ih,iw,ci] = sum_{kh,kw,co} dy[n,oh,ow,co] * w[kh,kw,ci,co]                     [K10]
B2C_API void b2c_conv2d_dgrad(const float* dy, const float* w, float* dx, int N, int H, int W, int Ci,
                              int Co, int KH, int KW, int sh, int sw, int ph, int pw, int OH, int OW) {
  const ConvShape s = make_shape(N, H, W, Ci, Co, KH, KW, sh, sw, ph, pw, OH, OW);
  if (KH * KW * Ci >= kGemmConvK) {
    // dcol[pix, K] = dy[pix, Co] . W^T, then scattered back onto the input pixels
    const int K = KH * KW * Ci;
    const int64_t pix = (int64_t)N * s.OH * s.OW;
    std::vector<float> dcol((size_t)pix * K);
    b2c_gemm(dy, Co, 1, w, 1, Co, dcol.data(), K, nullptr, (int)pix, K, Co, 0, 0);
    col2im(dcol.data(), dx, s);
    return;
  }
  parallel_for((int64_t)N * H, 4, [&](int64_t b, int64_t e) {
    for (int64_t r = b; r < e; ++r) {
      const int n = r / H, ih = r % H;
      for (int iw = 0; iw < W; ++iw) {
        float* dxp = dx + (((int64_t)n * H + ih) * W + iw) * Ci;
        for (int ci = 0; ci < Ci; ++ci) dxp[ci] = 0.f;
        for (int kh = 0; kh < KH; ++kh) {
          const int th = ih + ph - kh;
          if (th < 0 || th % sh) continue;
          const int oh = th / sh;
          if (oh >= s.OH) continue;
          for (int kw = 0; kw < KW; ++kw) {
            const int tw = iw + pw - kw;
            if (tw < 0 || tw % sw) continue;
            const int ow = tw / sw;
            if (ow >= s.OW) continue;
            const float* dp = dy + (((int64_t)n * s.OH + oh) * s.OW + ow) * Co;
            const float* wp = w + (int64_t)(kh * KW + kw) * Ci * Co;
            for (int ci = 0; ci < Ci; ++ci) {
              const float* wr = wp + (int64_t)ci * Co;
              float a = 0.f;
              for (int c = 0; c < Co; ++c) a += dp[c] * wr[c];
              dxp[ci] += a;
            }
          }
        }
      }
    }
  });
}

// 2x2/stride-2 max pool with argmax code (0..3 = dy*2+dx; first maximum wins)        [K2]
// Inner loops run over the contiguous channel dimension with select-style updates (no branches), which
// the compiler vectorises.
B2C_API void b2c_maxpool2_fwd(const float* x, float* y, uint8_t* code, int N, int H, int W, int C) {
  const int PH = H / 2, PW = W / 2;
  parallel_for((int64_t)N * PH, 4, [&](int64_t b, int64_t e) {
    for (int64_t r = b; r < e; ++r) {
      const int n = r / PH, ph = r % PH;
      for (int pw = 0; pw < PW; ++pw) {
        const float* __restrict__ p00 = x + (((int64_t)n * H + 2 * ph) * W + 2 * pw) * C;
        const float* __restrict__ p01 = p00 + C;
        const float* __restrict__ p10 = p00 + (int64_t)W * C;
        const float* __restrict__ p11 = p10 + C;
        float* __restrict__ yo = y + (((int64_t)n * PH + ph) * PW + pw) * C;
        uint8_t* __restrict__ co = code + (((int64_t)n * PH + ph) * PW + pw) * C;
        for (int c = 0; c < C; ++c) {
          float best = p00[c];
          int arg = 0;
          const float v1 = p01[c], v2 = p10[c], v3 = p11[c];
          arg = v1 > best ? 1 : arg, best = v1 > best ? v1 : best;
          arg = v2 > best ? 2 : arg, best = v2 > best ? v2 : best;
          arg = v3 > best ? 3 : arg, best = v3 > best ? v3 : best;
          yo[c] = best;
          co[c] = (uint8_t)arg;
        }
      }
    }
  });
}

// Pool backward, optionally fused with the ReLU backward of the producing layer.     [K8]
B2C_API void b2c_maxpool2_bwd(const float* dy, const uint8_t* code, const float* y_pooled, float* dx,
                              int N, int H, int W, int C, int act_relu) {
  const int PH = H / 2, PW = W / 2;
  parallel_for((int64_t)N * H, 4, [&](int64_t b, int64_t e) {
    for (int64_t r = b; r < e; ++r) {
      const int n = r / H, hh = r % H;
      const int ph = hh >> 1;
      for (int w = 0; w < W; ++w) {
        float* __restrict__ dxo = dx + (((int64_t)n * H + hh) * W + w) * C;
        const int pw = w >> 1;
        if (ph >= PH || pw >= PW) {          // odd trailing row / column: outside every window
          for (int c = 0; c < C; ++c) dxo[c] = 0.f;
          continue;
        }
        const int64_t o = (((int64_t)n * PH + ph) * PW + pw) * C;
        const float* __restrict__ g = dy + o;
        const uint8_t* __restrict__ cd = code + o;
        const float* __restrict__ yp = y_pooled + o;
        const int me = (hh & 1) * 2 + (w & 1);
        if (act_relu) {
          for (int c = 0; c < C; ++c) dxo[c] = (cd[c] == me && yp[c] > 0.f) ? g[c] : 0.f;
        } else {
          for (int c = 0; c < C; ++c) dxo[c] = cd[c] == me ? g[c] : 0.f;
        }
      }
    }
  });
}

// D[m,:] (+)= act(A[m,:] . B + bias), arbitrary strides so one routine serves forward
// (x.W), dgrad (dy.W^T) and wgrad (x^T.dy).                                      [K3/K4/K6/K7]
// B is first packed into zero-padded panels of kNR columns ([panel][k][kNR], contiguous), then a
// kMR x kNR register tile walks K: two 8-float vectors of B and one broadcast of A per FMA row, which the
// compiler turns into AVX2 FMAs without reassociating any sum (every output is accumulated in k order, so
// the result does not depend on the thread count and is identical on every worker).
namespace {
constexpr int kMR = 4, kNR = 16;

// GCC vector extension: 8 floats; the unaligned variant is for loads from the packed panels.
typedef float v8 __attribute__((vector_size(32)));
typedef float v8u __attribute__((vector_size(32), aligned(4)));

inline void gemm_tile(const float* A, int64_t sam, int64_t sak, const float* panel, int K, int rows,
                      float (&c)[kMR][kNR]) {
  const float* a0 = A;
  const float* a1 = A + (rows > 1 ? 1 : 0) * sam;
  const float* a2 = A + (rows > 2 ? 2 : 0) * sam;
  const float* a3 = A + (rows > 3 ? 3 : 0) * sam;
  v8 c00 = *reinterpret_cast<const v8u*>(&c[0][0]), c01 = *reinterpret_cast<const v8u*>(&c[0][8]);
  v8 c10 = *reinterpret_cast<const v8u*>(&c[1][0]), c11 = *reinterpret_cast<const v8u*>(&c[1][8]);
  v8 c20 = *reinterpret_cast<const v8u*>(&c[2][0]), c21 = *reinterpret_cast<const v8u*>(&c[2][8]);
  v8 c30 = *reinterpret_cast<const v8u*>(&c[3][0]), c31 = *reinterpret_cast<const v8u*>(&c[3][8]);
  for (int k = 0; k < K; ++k) {
    const v8 b0 = *reinterpret_cast<const v8u*>(panel + (int64_t)k * kNR);
    const v8 b1 = *reinterpret_cast<const v8u*>(panel + (int64_t)k * kNR + 8);
    const float s0 = a0[k * sak], s1 = a1[k * sak], s2 = a2[k * sak], s3 = a3[k * sak];
    const v8 v0 = {s0, s0, s0, s0, s0, s0, s0, s0}, v1 = {s1, s1, s1, s1, s1, s1, s1, s1};
    const v8 v2 = {s2, s2, s2, s2, s2, s2, s2, s2}, v3 = {s3, s3, s3, s3, s3, s3, s3, s3};
    c00 += v0 * b0, c01 += v0 * b1;
    c10 += v1 * b0, c11 += v1 * b1;
    c20 += v2 * b0, c21 += v2 * b1;
    c30 += v3 * b0, c31 += v3 * b1;
  }
  *reinterpret_cast<v8u*>(&c[0][0]) = c00, *reinterpret_cast<v8u*>(&c[0][8]) = c01;
  *reinterpret_cast<v8u*>(&c[1][0]) = c10, *reinterpret_cast<v8u*>(&c[1][8]) = c11;
  *reinterpret_cast<v8u*>(&c[2][0]) = c20, *reinterpret_cast<v8u*>(&c[2][8]) = c21;
  *reinterpret_cast<v8u*>(&c[3][0]) = c30, *reinterpret_cast<v8u*>(&c[3][8]) = c31;
}
}  // namespace

B2C_API void b2c_gemm(const float* A, int64_t sam, int64_t sak, const float* B, int64_t sbk,
                      int64_t sbn, float* D, int64_t ldd, const float* bias, int M, int N, int K,
                      int relu, int accumulate) {
  const int panels = (N + kNR - 1) / kNR;
  std::vector<float> packed((size_t)panels * K * kNR);
  parallel_for(panels, 1, [&](int64_t p0, int64_t p1) {
    for (int64_t p = p0; p < p1; ++p) {
      float* dst = packed.data() + p * (int64_t)K * kNR;
      const int n0 = (int)p * kNR, nr = std::min(kNR, N - n0);
      for (int k = 0; k < K; ++k) {
        const float* src = B + k * sbk + (int64_t)n0 * sbn;
        for (int j = 0; j < nr; ++j) dst[k * kNR + j] = src[j * sbn];
        for (int j = nr; j < kNR; ++j) dst[k * kNR + j] = 0.f;
      }
    }
  });
  const int mblocks = (M + kMR - 1) / kMR;
  // one task = one (panel, row block); panels outermost so a thread keeps its B panel in cache
  parallel_for((int64_t)panels * mblocks, 8, [&](int64_t t0, int64_t t1) {
    for (int64_t t = t0; t < t1; ++t) {
      const int p = (int)(t / mblocks), mb = (int)(t % mblocks);
      const int m0 = mb * kMR, rows = std::min(kMR, M - m0);
      const int n0 = p * kNR, nr = std::min(kNR, N - n0);
      float c[kMR][kNR];
      for (int r = 0; r < kMR; ++r)
        for (int j = 0; j < kNR; ++j) c[r][j] = (bias != nullptr && j < nr) ? bias[n0 + j] : 0.f;
      gemm_tile(A + (int64_t)m0 * sam, sam, sak, packed.data() + (int64_t)p * K * kNR, K, rows, c);
      for (int r = 0; r < rows; ++r) {
        float* dr = D + (int64_t)(m0 + r) * ldd + n0;
        for (int j = 0; j < nr; ++j) {
          const float v = relu ? std::max(c[r][j], 0.f) : c[r][j];
          dr[j] = accumulate ? dr[j] + v : v;
        }
      }
    }
  });
}

// Softmax cross-entropy on logits: dz = (softmax - onehot) * grad_scale ;
// metrics[0] += sum loss, metrics[1] += #correct, metrics[2] += #rows.                 [K5]
B2C_API void b2c_softmax_xent(const float* z, const int32_t* labels, float* dz, float* metrics, int B,
                              int C, float grad_scale) {
  double loss = 0.0;
  int correct = 0;
  for (int r = 0; r < B; ++r) {
    const float* zr = z + (int64_t)r * C;
    float mx = zr[0];
    int arg = 0;
    for (int c = 1; c < C; ++c)
      if (zr[c] > mx) mx = zr[c], arg = c;
    float se = 0.f;
    for (int c = 0; c < C; ++c) se += std::exp(zr[c] - mx);
    const int y = labels[r];
    loss += (double)(mx + std::log(se) - zr[y]);
    correct += arg == y;
    if (dz != nullptr) {
      const float inv = 1.f / se;
      for (int c = 0; c < C; ++c)
        dz[(int64_t)r * C + c] = (std::exp(zr[c] - mx) * inv - (c == y ? 1.f : 0.f)) * grad_scale;
    }
  }
  if (metrics != nullptr) {
    metrics[0] += (float)loss;
    metrics[1] += (float)correct;
    metrics[2] += (float)B;
  }
}

B2C_API void b2c_relu_bwd(const float* dy, const float* y, float* dx, int64_t n) {
  for (int64_t i = 0; i < n; ++i) dx[i] = y[i] > 0.f ? dy[i] : 0.f;
}

B2C_API void b2c_colsum(const float* dy, float* db, int M, int N) {
  for (int64_t m = 0; m < M; ++m)
    for (int n = 0; n < N; ++n) db[n] += dy[m * N + n];
}

// w -= lr * g over the parameters, tail slots (metrics) accumulate; g is zeroed.      [K13]
B2C_API void b2c_sgd_apply(float* w, float* g, int64_t n_params, int64_t n_total, float lr) {
  for (int64_t i = 0; i < n_params; ++i) w[i] -= lr * g[i];
  for (int64_t i = n_params; i < n_total; ++i) w[i] += g[i];
  std::memset(g, 0, sizeof(float) * n_total);
}
