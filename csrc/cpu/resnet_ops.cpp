// resnet_ops.cpp — C++ CPU counterparts of csrc/cuda/resnet_kernels.cu (libb200dist_cpu.so):
// BatchNormalization (training / inference, fused ReLU, moving-statistics pseudo-gradients),
// residual add (+ReLU / masked), general max pooling, global average pooling.  They are the fp32
// oracle for the CUDA kernels and let residual networks train on the CPU backend.  None of these
// layers appears in the reference README (its model is README.md:292-298).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <vector>

#define B2C_API extern "C" __attribute__((visibility("default")))

// Training: batch statistics over the M rows of x [M, C]; saved = mean | rstd for backward.
// Moving statistics: with g_mean / g_var given, the pseudo-gradient (1-momentum)*(m - batch)*inv_lr is
// stored there (the optimizer's w -= lr*g then performs the moving-average update, averaged over
// replicas by the gradient all-reduce); otherwise they are updated in place.  Inference: normalise
// with the moving statistics and touch nothing else.  residual (optional, like x): y = act(bn(x) + residual).
B2C_API void b2c_bn_fwd(const float* x, const float* gamma, const float* beta, float* y, float* saved,
                        float* moving_mean, float* moving_var, float* g_mean, float* g_var, int64_t M, int C,
                        float eps, float momentum, float inv_lr, int relu, int inference, const float* residual) {
  std::vector<double> sum(C, 0.0), sq(C, 0.0);
  std::vector<float> sc(C), sh(C);
  if (!inference) {
    for (int64_t r = 0; r < M; ++r)
      for (int c = 0; c < C; ++c) {
        const double v = x[r * C + c];
        sum[c] += v, sq[c] += v * v;
      }
  }
  for (int c = 0; c < C; ++c) {
    float mean, var;
    if (inference) {
      mean = moving_mean[c], var = moving_var[c];
    } else {
      mean = (float)(sum[c] / (double)M);
      var = (float)std::max(sq[c] / (double)M - (sum[c] / (double)M) * (sum[c] / (double)M), 0.0);
    }
    const float rstd = 1.f / std::sqrt(var + eps);
    sc[c] = gamma[c] * rstd;
    sh[c] = beta[c] - mean * sc[c];
    if (!inference) {
      saved[c] = mean, saved[C + c] = rstd;
      const float unbiased = M > 1 ? var * (float)M / (float)(M - 1) : var;
      if (g_mean != nullptr) {
        g_mean[c] = (1.f - momentum) * (moving_mean[c] - mean) * inv_lr;
        g_var[c] = (1.f - momentum) * (moving_var[c] - unbiased) * inv_lr;
      } else if (moving_mean != nullptr) {
        moving_mean[c] = momentum * moving_mean[c] + (1.f - momentum) * mean;
        moving_var[c] = momentum * moving_var[c] + (1.f - momentum) * unbiased;
      }
    }
  }
  for (int64_t r = 0; r < M; ++r)
    for (int c = 0; c < C; ++c) {
      const float v = x[r * C + c] * sc[c] + sh[c] + (residual ? residual[r * C + c] : 0.f);
      y[r * C + c] = relu ? std::max(v, 0.f) : v;
    }
}

// dgamma += sum g*xhat, dbeta += sum g, dx = gamma*rstd*(g - mean(g) - xhat*mean(g*xhat)); g = dy
// masked by y > 0 when relu.  dx may be NULL.
B2C_API void b2c_bn_bwd(const float* x, const float* y, const float* dy, const float* gamma, const float* saved,
                        float* dx, float* dgamma, float* dbeta, int64_t M, int C, int relu) {
  std::vector<double> sg(C, 0.0), sgx(C, 0.0);
  for (int64_t r = 0; r < M; ++r)
    for (int c = 0; c < C; ++c) {
      const int64_t o = r * C + c;
      const float g = (relu && !(y[o] > 0.f)) ? 0.f : dy[o];
      sg[c] += g;
      sgx[c] += (double)g * (x[o] - saved[c]) * saved[C + c];
    }
  for (int c = 0; c < C; ++c) dbeta[c] += (float)sg[c], dgamma[c] += (float)sgx[c];
  if (dx == nullptr) return;
  const double inv_m = 1.0 / (double)M;
  for (int64_t r = 0; r < M; ++r)
    for (int c = 0; c < C; ++c) {
      const int64_t o = r * C + c;
      const float g = (relu && !(y[o] > 0.f)) ? 0.f : dy[o];
      const float xhat = (x[o] - saved[c]) * saved[C + c];
      dx[o] = gamma[c] * saved[C + c] * (g - (float)(sg[c] * inv_m) - xhat * (float)(sgx[c] * inv_m));
    }
}

// mode 0: out = a + b; 1: relu(a + b); 2: (a + b) where mask > 0 else 0.  b may be NULL.
B2C_API void b2c_add_act(const float* a, const float* b, const float* mask, float* out, int64_t n, int mode) {
  for (int64_t i = 0; i < n; ++i) {
    float v = a[i] + (b ? b[i] : 0.f);
    if (mode == 1) v = std::max(v, 0.f);
    if (mode == 2 && !(mask[i] > 0.f)) v = 0.f;
    out[i] = v;
  }
}

// General max pooling: window k, stride s, top/left padding p, explicit output size.
B2C_API void b2c_maxpool_fwd(const float* x, float* y, uint8_t* code, int N, int H, int W, int C, int k, int s,
                             int p, int OH, int OW) {
  for (int n = 0; n < N; ++n)
    for (int oh = 0; oh < OH; ++oh)
      for (int ow = 0; ow < OW; ++ow)
        for (int c = 0; c < C; ++c) {
          float best = -INFINITY;
          int arg = 0;
          for (int kh = 0; kh < k; ++kh) {
            const int ih = oh * s - p + kh;
            if (ih < 0 || ih >= H) continue;
            for (int kw = 0; kw < k; ++kw) {
              const int iw = ow * s - p + kw;
              if (iw < 0 || iw >= W) continue;
              const float v = x[(((int64_t)n * H + ih) * W + iw) * C + c];
              if (v > best) best = v, arg = kh * k + kw;
            }
          }
          const int64_t o = (((int64_t)n * OH + oh) * OW + ow) * C + c;
          y[o] = best;
          code[o] = (uint8_t)arg;
        }
}

B2C_API void b2c_maxpool_bwd(const float* dy, const uint8_t* code, float* dx, int N, int H, int W, int C, int k,
                             int s, int p, int OH, int OW) {
  const int64_t total = (int64_t)N * H * W * C;
  for (int64_t i = 0; i < total; ++i) dx[i] = 0.f;
  for (int n = 0; n < N; ++n)
    for (int oh = 0; oh < OH; ++oh)
      for (int ow = 0; ow < OW; ++ow)
        for (int c = 0; c < C; ++c) {
          const int64_t o = (((int64_t)n * OH + oh) * OW + ow) * C + c;
          const int kh = code[o] / k, kw = code[o] % k;
          const int ih = oh * s - p + kh, iw = ow * s - p + kw;
          if (ih >= 0 && ih < H && iw >= 0 && iw < W) dx[(((int64_t)n * H + ih) * W + iw) * C + c] += dy[o];
        }
}

// Global average pooling and its backward (dx zeroed where mask <= 0; mask may be NULL).
B2C_API void b2c_gap_fwd(const float* x, float* y, int N, int HW, int C) {
  for (int n = 0; n < N; ++n)
    for (int c = 0; c < C; ++c) {
      double s = 0.0;
      for (int p = 0; p < HW; ++p) s += x[((int64_t)n * HW + p) * C + c];
      y[(int64_t)n * C + c] = (float)(s / HW);
    }
}

B2C_API void b2c_gap_bwd(const float* dy, const float* mask, float* dx, int N, int HW, int C) {
  const float inv = 1.f / (float)HW;
  for (int n = 0; n < N; ++n)
    for (int p = 0; p < HW; ++p)
      for (int c = 0; c < C; ++c) {
        const int64_t o = ((int64_t)n * HW + p) * C + c;
        dx[o] = (mask != nullptr && !(mask[o] > 0.f)) ? 0.f : dy[(int64_t)n * C + c] * inv;
      }
}
