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
#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>
#include <thread>
#include <vector>

#define B2C_API extern "C" __attribute__((visibility("default")))

namespace {

int g_threads = 0;

int num_threads() {
  if (g_threads > 0) return g_threads;
  unsigned hw = std::thread::hardware_concurrency();
  return static_cast<int>(std::max(1u, std::min(hw, 8u)));
}

// Static-partition parallel for over [0, n); falls back to inline when the range is small.
void parallel_for(int64_t n, int64_t grain, const std::function<void(int64_t, int64_t)>& fn) {
  const int nt = static_cast<int>(std::min<int64_t>(num_threads(), (n + grain - 1) / grain));
  if (nt <= 1) {
    fn(0, n);
    return;
  }
  std::vector<std::thread> th;
  th.reserve(nt - 1);
  const int64_t chunk = (n + nt - 1) / nt;
  for (int t = 1; t < nt; ++t) {
    const int64_t b = t * chunk, e = std::min(n, b + chunk);
    if (b < e) th.emplace_back([=, &fn] { fn(b, e); });
  }
  fn(0, std::min(n, chunk));
  for (auto& t : th) t.join();
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

// y[n,oh,ow,:] = act(b + sum_{kh,kw,ci} x[n,ih,iw,ci] * w[kh,kw,ci,:])               [K1]
B2C_API void b2c_conv2d_fwd(const float* x, const float* w, const float* bias, float* y, int N, int H,
                            int W, int Ci, int Co, int KH, int KW, int sh, int sw, int ph, int pw,
                            int OH, int OW, int relu) {
  const ConvShape s = make_shape(N, H, W, Ci, Co, KH, KW, sh, sw, ph, pw, OH, OW);
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
  // Parallel over filter taps (kh, kw, ci): each owns a disjoint dW row, so no reduction.
  parallel_for((int64_t)KH * KW * Ci, 1, [&](int64_t b, int64_t e) {
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
B2C_API void b2c_maxpool2_fwd(const float* x, float* y, uint8_t* code, int N, int H, int W, int C) {
  const int PH = H / 2, PW = W / 2;
  parallel_for((int64_t)N * PH, 8, [&](int64_t b, int64_t e) {
    for (int64_t r = b; r < e; ++r) {
      const int n = r / PH, ph = r % PH;
      for (int pw = 0; pw < PW; ++pw)
        for (int c = 0; c < C; ++c) {
          const float* p = x + (((int64_t)n * H + 2 * ph) * W + 2 * pw) * C + c;
          float best = p[0];
          int arg = 0;
          if (p[C] > best) best = p[C], arg = 1;
          if (p[(int64_t)W * C] > best) best = p[(int64_t)W * C], arg = 2;
          if (p[(int64_t)W * C + C] > best) best = p[(int64_t)W * C + C], arg = 3;
          const int64_t o = (((int64_t)n * PH + ph) * PW + pw) * C + c;
          y[o] = best;
          code[o] = (uint8_t)arg;
        }
    }
  });
}

// Pool backward, optionally fused with the ReLU backward of the producing layer.     [K8]
B2C_API void b2c_maxpool2_bwd(const float* dy, const uint8_t* code, const float* y_pooled, float* dx,
                              int N, int H, int W, int C, int act_relu) {
  const int PH = H / 2, PW = W / 2;
  parallel_for((int64_t)N * H, 8, [&](int64_t b, int64_t e) {
    for (int64_t r = b; r < e; ++r) {
      const int n = r / H, hh = r % H;
      for (int w = 0; w < W; ++w)
        for (int c = 0; c < C; ++c) {
          const int ph = hh >> 1, pw = w >> 1;
          float g = 0.f;
          if (ph < PH && pw < PW) {
            const int64_t o = (((int64_t)n * PH + ph) * PW + pw) * C + c;
            if (code[o] == ((hh & 1) * 2 + (w & 1)) && !(act_relu && !(y_pooled[o] > 0.f))) g = dy[o];
          }
          dx[(((int64_t)n * H + hh) * W + w) * C + c] = g;
        }
    }
  });
}

// D[m,:] (+)= act(A[m,:] . B + bias), arbitrary strides so one routine serves forward
// (x.W), dgrad (dy.W^T) and wgrad (x^T.dy).                                      [K3/K4/K6/K7]
B2C_API void b2c_gemm(const float* A, int64_t sam, int64_t sak, const float* B, int64_t sbk,
                      int64_t sbn, float* D, int64_t ldd, const float* bias, int M, int N, int K,
                      int relu, int accumulate) {
  parallel_for(M, 4, [&](int64_t b, int64_t e) {
    std::vector<float> acc(N);
    for (int64_t m = b; m < e; ++m) {
      for (int n = 0; n < N; ++n) acc[n] = bias ? bias[n] : 0.f;
      if (sbn == 1) {
        for (int k = 0; k < K; ++k) {
          const float av = A[m * sam + k * sak];
          const float* br = B + k * sbk;
          for (int n = 0; n < N; ++n) acc[n] += av * br[n];
        }
      } else {
        for (int n = 0; n < N; ++n) {
          float a = 0.f;
          const float* bc = B + n * sbn;
          for (int k = 0; k < K; ++k) a += A[m * sam + k * sak] * bc[k * sbk];
          acc[n] += a;
        }
      }
      float* dr = D + m * ldd;
      for (int n = 0; n < N; ++n) {
        const float v = relu ? std::max(acc[n], 0.f) : acc[n];
        dr[n] = accumulate ? dr[n] + v : v;
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
