// resnet_kernels.cu — the memory-bound kernels of residual networks (BASELINE.json config 4),
// templated on the activation storage type: fp32 or bf16 (fp32 arithmetic either way).
//
//   * BatchNormalization: batch statistics, normalise (+ReLU) fused with the saved / moving
//     statistics bookkeeping, inference mode on the moving statistics, and the two-pass backward
//   * residual add (+ReLU), masked add for the backward of a block whose input carries a ReLU
//   * general max pooling (overlapping windows) forward / backward, global average pooling
//   * im2col for small-channel convolutions, dtype casts, bias-gradient column sums
//
// None of these layers appears in the reference README (its model is README.md:292-298); they
// exist because the target configurations name ResNet-50.  All tensors are NHWC, viewed as
// [M = N*H*W rows][C channels]; every access is a 16-byte vector (4 fp32 / 8 bf16 channels).
//
// Thread mapping shared by the per-channel kernels: a CTA of 256 threads covers `lanes_c`
// channel vectors (a power of two <= 32) x 256/lanes_c row groups, so narrow tensors (C = 64)
// still issue full 512-byte warp accesses; a thread keeps its channels for the whole kernel, which
// is what lets per-channel constants (mean, rstd, gamma...) live in registers.
#include <cuda_bf16.h>

#include "host_common.h"
#include "ptx.cuh"

namespace b2d {

template <typename T>
struct VecIO;

template <>
struct VecIO<float> {
  static constexpr int N = 4;
  static __device__ __forceinline__ void load(const float* p, float (&v)[4]) {
    const float4 t = *reinterpret_cast<const float4*>(p);
    v[0] = t.x, v[1] = t.y, v[2] = t.z, v[3] = t.w;
  }
  static __device__ __forceinline__ void store(float* p, const float (&v)[4]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  }
};

template <>
struct VecIO<__nv_bfloat16> {
  static constexpr int N = 8;
  static __device__ __forceinline__ void load(const __nv_bfloat16* p, float (&v)[8]) {
    const uint4 t = *reinterpret_cast<const uint4*>(p);
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      v[2 * k] = __uint_as_float(w[k] << 16);
      v[2 * k + 1] = __uint_as_float(w[k] & 0xffff0000u);
    }
  }
  static __device__ __forceinline__ void store(__nv_bfloat16* p, const float (&v)[8]) {
    uint32_t w[4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
      asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(w[k]) : "f"(v[2 * k + 1]), "f"(v[2 * k]));
    *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
  }
};

__device__ __forceinline__ void store_scalar(float* p, float v) { *p = v; }
__device__ __forceinline__ void store_scalar(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }
__device__ __forceinline__ float load_scalar(const float* p) { return *p; }
__device__ __forceinline__ float load_scalar(const __nv_bfloat16* p) { return __bfloat162float(*p); }

// One 16-byte reduction per four channels: narrow tensors (C = 64) funnel every CTA's partial sums
// into a handful of L2 lines, so the number of atomic operations is what bounds the tail.
__device__ __forceinline__ void red_add_v4(float* p, const float* v) {
  asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v[0]), "f"(v[1]), "f"(v[2]),
               "f"(v[3])
               : "memory");
}

struct ChanMap {
  int cl, rg, nrg, col;
  bool ok;
};

template <int V>
__device__ __forceinline__ ChanMap chan_map(int C, int lanes_c) {
  ChanMap m;
  m.cl = threadIdx.x & (lanes_c - 1);
  m.rg = threadIdx.x / lanes_c;
  m.nrg = blockDim.x / lanes_c;
  m.col = (blockIdx.x * lanes_c + m.cl) * V;
  m.ok = m.col < C;
  return m;
}

// Sums s[] (and q[]) over all threads of the CTA that share a channel vector; the result is
// valid in warp 0, lanes < lanes_c.
template <int V, bool TWO>
__device__ __forceinline__ void cta_channel_reduce(float (&s)[V], float (&q)[V], int lanes_c) {
  __shared__ float ps[8][32][V];
  __shared__ float pq[TWO ? 8 : 1][32][V];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int off = lanes_c; off < 32; off <<= 1) {
#pragma unroll
    for (int k = 0; k < V; ++k) {
      s[k] += __shfl_xor_sync(0xffffffffu, s[k], off);
      if (TWO) q[k] += __shfl_xor_sync(0xffffffffu, q[k], off);
    }
  }
  if (lane < lanes_c) {
#pragma unroll
    for (int k = 0; k < V; ++k) {
      ps[warp][lane][k] = s[k];
      if (TWO) pq[warp][lane][k] = q[k];
    }
  }
  __syncthreads();
  if (warp == 0 && lane < lanes_c) {
#pragma unroll
    for (int k = 0; k < V; ++k) {
      float a = 0.f, b = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) {
        a += ps[w][lane][k];
        if (TWO) b += pq[w][lane][k];
      }
      s[k] = a;
      if (TWO) q[k] = b;
    }
  }
}

// ---------------------------------------------------------------------------------------
// Column sums (and sums of squares): out[0:C] += sum_r x[r,:], out[C:2C] += sum_r x[r,:]^2.
// SQ = false is the bias gradient of a convolution (db += column sums of dy).
// ---------------------------------------------------------------------------------------
template <typename T, bool SQ>
__global__ void __launch_bounds__(256) col_stats_kernel(const T* __restrict__ x, float* __restrict__ out, int M,
                                                         int C, int rows_per_cta, int lanes_c) {
  constexpr int V = VecIO<T>::N;
  pdl_wait();
  pdl_trigger();
  const ChanMap m = chan_map<V>(C, lanes_c);
  const int r0 = blockIdx.y * rows_per_cta, r1 = min(M, r0 + rows_per_cta);
  float s[V], q[V];
#pragma unroll
  for (int k = 0; k < V; ++k) s[k] = 0.f, q[k] = 0.f;
  if (m.ok) {
#pragma unroll 8
    for (int r = r0 + m.rg; r < r1; r += m.nrg) {
      float v[V];
      VecIO<T>::load(x + (size_t)r * C + m.col, v);
#pragma unroll
      for (int k = 0; k < V; ++k) {
        s[k] += v[k];
        if (SQ) q[k] += v[k] * v[k];
      }
    }
  }
  cta_channel_reduce<V, SQ>(s, q, lanes_c);
  if (threadIdx.x < lanes_c && m.ok) {
#pragma unroll
    for (int k = 0; k < V; k += 4) {
      red_add_v4(out + m.col + k, s + k);
      if (SQ) red_add_v4(out + C + m.col + k, q + k);
    }
  }
}

struct BnArgs {
  const void* x;
  void* y;             // forward output; backward: the ReLU mask source when relu != 0
  const void* dy;
  void* dx;
  const void* residual;  // forward: y = act(bn(x) + residual); backward then applies no ReLU mask
  const float *gamma, *beta;
  float *moving_mean, *moving_var;            // read in inference mode; updated in place when no *_grad
  float *moving_mean_grad, *moving_var_grad;  // training: pseudo-gradients (see b200dist.h)
  float *stats, *saved, *red;
  float *dgamma, *dbeta;
  int M, C, relu, inference;
  float eps, momentum, inv_lr;
  int rows_per_cta, lanes_c;
};

// y = act(gamma * (x - mean) * rstd + beta).  Training: mean / var from the batch sums in
// `stats`; the CTA row 0 also publishes mean|rstd to `saved`, advances the moving statistics and
// clears `red` for this step's backward pass.  Inference: mean / var are the moving statistics.
template <typename T>
__global__ void __launch_bounds__(256) bn_apply_kernel(const BnArgs a) {
  constexpr int V = VecIO<T>::N;
  pdl_wait();
  pdl_trigger();
  const int C = a.C, M = a.M;
  const ChanMap m = chan_map<V>(C, a.lanes_c);
  if (!m.ok) return;
  const float inv_m = 1.f / (float)M;
  float sc[V], sh[V];
#pragma unroll
  for (int k = 0; k < V; ++k) {
    const int c = m.col + k;
    float mean, var;
    if (a.inference) {
      mean = a.moving_mean[c], var = a.moving_var[c];
    } else {
      mean = a.stats[c] * inv_m;
      var = fmaxf(a.stats[C + c] * inv_m - mean * mean, 0.f);
    }
    const float rstd = rsqrtf(var + a.eps);
    sc[k] = a.gamma[c] * rstd;
    sh[k] = a.beta[c] - mean * sc[k];
    if (!a.inference && blockIdx.y == 0 && m.rg == 0) {
      a.saved[c] = mean, a.saved[C + c] = rstd;
      a.red[c] = 0.f, a.red[C + c] = 0.f;
      const float unbiased = M > 1 ? var * (float)M / (float)(M - 1) : var;
      if (a.moving_mean_grad != nullptr) {
        // w -= lr * g in the optimizer turns this into momentum*w + (1-momentum)*batch, averaged
        // over replicas by the gradient all-reduce
        a.moving_mean_grad[c] = (1.f - a.momentum) * (a.moving_mean[c] - mean) * a.inv_lr;
        a.moving_var_grad[c] = (1.f - a.momentum) * (a.moving_var[c] - unbiased) * a.inv_lr;
      } else if (a.moving_mean != nullptr) {
        a.moving_mean[c] = a.momentum * a.moving_mean[c] + (1.f - a.momentum) * mean;
        a.moving_var[c] = a.momentum * a.moving_var[c] + (1.f - a.momentum) * unbiased;
      }
    }
  }
  const T* x = static_cast<const T*>(a.x);
  const T* res = static_cast<const T*>(a.residual);
  T* y = static_cast<T*>(a.y);
  const int r0 = blockIdx.y * a.rows_per_cta, r1 = min(M, r0 + a.rows_per_cta);
#pragma unroll 4
  for (int r = r0 + m.rg; r < r1; r += m.nrg) {
    float v[V];
    VecIO<T>::load(x + (size_t)r * C + m.col, v);
    if (res != nullptr) {
      float rv[V];
      VecIO<T>::load(res + (size_t)r * C + m.col, rv);
#pragma unroll
      for (int k = 0; k < V; ++k) v[k] = fmaf(v[k], sc[k], sh[k]) + rv[k];
    } else {
#pragma unroll
      for (int k = 0; k < V; ++k) v[k] = fmaf(v[k], sc[k], sh[k]);
    }
    if (a.relu) {
#pragma unroll
      for (int k = 0; k < V; ++k) v[k] = fmaxf(v[k], 0.f);
    }
    VecIO<T>::store(y + (size_t)r * C + m.col, v);
  }
}

// Backward pass 1: red[0:C] += sum g, red[C:2C] += sum g * xhat, with g = dy * (y > 0 if relu).
template <typename T>
__global__ void __launch_bounds__(256) bn_bwd_reduce_kernel(const BnArgs a) {
  constexpr int V = VecIO<T>::N;
  pdl_wait();
  pdl_trigger();
  const int C = a.C, M = a.M;
  const ChanMap m = chan_map<V>(C, a.lanes_c);
  const T* x = static_cast<const T*>(a.x);
  const T* dy = static_cast<const T*>(a.dy);
  const int r0 = blockIdx.y * a.rows_per_cta, r1 = min(M, r0 + a.rows_per_cta);
  float s[V], q[V];
#pragma unroll
  for (int k = 0; k < V; ++k) s[k] = 0.f, q[k] = 0.f;
  if (m.ok) {
    // The ReLU mask (y > 0) is recomputed from x with the forward pass' own scale / shift instead of
    // reading y back: same fp32 expression, and rounding y to bf16 cannot change its sign.
    float mean[V], rstd[V], sc[V], sh[V];
#pragma unroll
    for (int k = 0; k < V; ++k) {
      mean[k] = a.saved[m.col + k], rstd[k] = a.saved[C + m.col + k];
      sc[k] = a.gamma[m.col + k] * rstd[k];
      sh[k] = a.beta[m.col + k] - mean[k] * sc[k];
    }
#pragma unroll 4
    for (int r = r0 + m.rg; r < r1; r += m.nrg) {
      const size_t o = (size_t)r * C + m.col;
      float g[V], xv[V];
      VecIO<T>::load(dy + o, g);
      VecIO<T>::load(x + o, xv);
      if (a.relu) {
#pragma unroll
        for (int k = 0; k < V; ++k) g[k] = fmaf(xv[k], sc[k], sh[k]) > 0.f ? g[k] : 0.f;
      }
#pragma unroll
      for (int k = 0; k < V; ++k) {
        s[k] += g[k];
        q[k] += g[k] * (xv[k] - mean[k]) * rstd[k];
      }
    }
  }
  cta_channel_reduce<V, true>(s, q, a.lanes_c);
  if (threadIdx.x < a.lanes_c && m.ok) {
#pragma unroll
    for (int k = 0; k < V; k += 4) {
      red_add_v4(a.red + m.col + k, s + k);
      red_add_v4(a.red + C + m.col + k, q + k);
    }
  }
}

// Backward pass 2: dx = gamma * rstd * (g - sum_g/M - xhat * sum_gxhat/M).  CTA row 0 also adds the
// sums into dgamma / dbeta and clears `stats` for the next forward pass (`red` is cleared there).
template <typename T>
__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(const BnArgs a) {
  constexpr int V = VecIO<T>::N;
  pdl_wait();
  pdl_trigger();
  const int C = a.C, M = a.M;
  const ChanMap m = chan_map<V>(C, a.lanes_c);
  if (!m.ok) return;
  const float inv_m = 1.f / (float)M;
  float mean[V], rstd[V], ga[V], sb[V], sq[V], sh[V];
#pragma unroll
  for (int k = 0; k < V; ++k) {
    const int c = m.col + k;
    mean[k] = a.saved[c], rstd[k] = a.saved[C + c];
    const float rs = a.red[c], rq = a.red[C + c];
    ga[k] = a.gamma[c] * rstd[k];
    sh[k] = a.beta[c] - mean[k] * ga[k];          // forward: y = x * ga + sh
    sb[k] = rs * inv_m, sq[k] = rq * inv_m;
    if (blockIdx.y == 0 && m.rg == 0) {
      a.dbeta[c] += rs, a.dgamma[c] += rq;
      a.stats[c] = 0.f, a.stats[C + c] = 0.f;
    }
  }
  if (a.dx == nullptr) return;
  const T* x = static_cast<const T*>(a.x);
  const T* dy = static_cast<const T*>(a.dy);
  T* dx = static_cast<T*>(a.dx);
  const int r0 = blockIdx.y * a.rows_per_cta, r1 = min(M, r0 + a.rows_per_cta);
#pragma unroll 4
  for (int r = r0 + m.rg; r < r1; r += m.nrg) {
    const size_t o = (size_t)r * C + m.col;
    float g[V], xv[V];
    VecIO<T>::load(dy + o, g);
    VecIO<T>::load(x + o, xv);
    if (a.relu) {     // mask recomputed from x (see bn_bwd_reduce_kernel)
#pragma unroll
      for (int k = 0; k < V; ++k) g[k] = fmaf(xv[k], ga[k], sh[k]) > 0.f ? g[k] : 0.f;
    }
#pragma unroll
    for (int k = 0; k < V; ++k) {
      const float xhat = (xv[k] - mean[k]) * rstd[k];
      g[k] = ga[k] * (g[k] - sb[k] - xhat * sq[k]);
    }
    VecIO<T>::store(dx + o, g);
  }
}

// mode 0: out = a + b      mode 1: out = relu(a + b)      mode 2: out = (a + b) * (m > 0)
// b may be NULL (treated as zero): mode 2 is then a stand-alone ReLU backward.
template <typename T>
__global__ void __launch_bounds__(256) add_act_kernel(const T* __restrict__ a, const T* __restrict__ b,
                                                       const T* __restrict__ msk, T* __restrict__ out, size_t nvec,
                                                       int mode) {
  constexpr int V = VecIO<T>::N;
  pdl_wait();
  pdl_trigger();
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nvec; i += (size_t)gridDim.x * blockDim.x) {
    float x[V], y[V];
    VecIO<T>::load(a + i * V, x);
    if (b != nullptr) {
      VecIO<T>::load(b + i * V, y);
#pragma unroll
      for (int k = 0; k < V; ++k) x[k] += y[k];
    }
    if (mode == 1) {
#pragma unroll
      for (int k = 0; k < V; ++k) x[k] = fmaxf(x[k], 0.f);
    } else if (mode == 2) {
      VecIO<T>::load(msk + i * V, y);
#pragma unroll
      for (int k = 0; k < V; ++k) x[k] = y[k] > 0.f ? x[k] : 0.f;
    }
    VecIO<T>::store(out + i * V, x);
  }
}

// ---------------------------------------------------------------------------------------
// General max pooling (window k, stride s, top/left padding p, explicit output size), NHWC.
// One thread = one output pixel x one channel vector.  code = index of the winning tap inside
// the window (kh * k + kw), one byte per element.
// ---------------------------------------------------------------------------------------
template <int V>
struct CodeVec;
template <>
struct CodeVec<4> {
  using type = uint32_t;
};
template <>
struct CodeVec<8> {
  using type = uint2;
};
__device__ __forceinline__ void unpack_code(uint32_t w, int (&c)[4]) {
#pragma unroll
  for (int k = 0; k < 4; ++k) c[k] = (w >> (8 * k)) & 0xff;
}
__device__ __forceinline__ void unpack_code(uint2 w, int (&c)[8]) {
#pragma unroll
  for (int k = 0; k < 4; ++k) c[k] = (w.x >> (8 * k)) & 0xff, c[4 + k] = (w.y >> (8 * k)) & 0xff;
}
__device__ __forceinline__ uint32_t pack_code(const int (&c)[4]) {
  return (uint32_t)c[0] | ((uint32_t)c[1] << 8) | ((uint32_t)c[2] << 16) | ((uint32_t)c[3] << 24);
}
__device__ __forceinline__ uint2 pack_code(const int (&c)[8]) {
  uint2 r;
  r.x = (uint32_t)c[0] | ((uint32_t)c[1] << 8) | ((uint32_t)c[2] << 16) | ((uint32_t)c[3] << 24);
  r.y = (uint32_t)c[4] | ((uint32_t)c[5] << 8) | ((uint32_t)c[6] << 16) | ((uint32_t)c[7] << 24);
  return r;
}

template <typename T>
__global__ void __launch_bounds__(256) maxpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y,
                                                           uint8_t* __restrict__ code, int N, int H, int W, int C,
                                                           int OH, int OW, int k, int s, int p) {
  constexpr int V = VecIO<T>::N;
  using CV = typename CodeVec<V>::type;
  pdl_wait();
  pdl_trigger();
  const int cv = C / V;
  const size_t total = (size_t)N * OH * OW * cv;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % cv) * V;
    size_t t = i / cv;
    const int ow = t % OW;
    t /= OW;
    const int oh = t % OH;
    const int n = (int)(t / OH);
    float best[V];
    int arg[V];
#pragma unroll
    for (int j = 0; j < V; ++j) best[j] = -INFINITY, arg[j] = 0;
    for (int kh = 0; kh < k; ++kh) {
      const int ih = oh * s - p + kh;
      if (ih < 0 || ih >= H) continue;
      for (int kw = 0; kw < k; ++kw) {
        const int iw = ow * s - p + kw;
        if (iw < 0 || iw >= W) continue;
        float v[V];
        VecIO<T>::load(x + (((size_t)n * H + ih) * W + iw) * C + c, v);
#pragma unroll
        for (int j = 0; j < V; ++j)
          if (v[j] > best[j]) best[j] = v[j], arg[j] = kh * k + kw;
      }
    }
    VecIO<T>::store(y + i * V, best);
    *reinterpret_cast<CV*>(code + i * V) = pack_code(arg);
  }
}

// Gather form of the backward pass (windows may overlap): dx[pixel] = sum of dy over the windows
// whose arg-max is this pixel.
template <typename T>
__global__ void __launch_bounds__(256) maxpool_bwd_kernel(const T* __restrict__ dy, const uint8_t* __restrict__ code,
                                                           T* __restrict__ dx, int N, int H, int W, int C, int OH,
                                                           int OW, int k, int s, int p) {
  constexpr int V = VecIO<T>::N;
  using CV = typename CodeVec<V>::type;
  pdl_wait();
  pdl_trigger();
  const int cv = C / V;
  const size_t total = (size_t)N * H * W * cv;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % cv) * V;
    size_t t = i / cv;
    const int iw = t % W;
    t /= W;
    const int ih = t % H;
    const int n = (int)(t / H);
    float g[V];
#pragma unroll
    for (int j = 0; j < V; ++j) g[j] = 0.f;
    for (int kh = 0; kh < k; ++kh) {
      const int th = ih + p - kh;
      if (th < 0 || th % s) continue;
      const int oh = th / s;
      if (oh >= OH) continue;
      for (int kw = 0; kw < k; ++kw) {
        const int tw = iw + p - kw;
        if (tw < 0 || tw % s) continue;
        const int ow = tw / s;
        if (ow >= OW) continue;
        const size_t o = (((size_t)n * OH + oh) * OW + ow) * C + c;
        int cd[V];
        unpack_code(*reinterpret_cast<const CV*>(code + o), cd);
        float v[V];
        VecIO<T>::load(dy + o, v);
#pragma unroll
        for (int j = 0; j < V; ++j)
          if (cd[j] == kh * k + kw) g[j] += v[j];
      }
    }
    VecIO<T>::store(dx + i * V, g);
  }
}

// Same gather with a compile-time window / stride: the (up to K*K) candidate windows of a pixel are enumerated
// without divisions, and ALL their (code, dy) pairs are fetched before any is used — one memory round trip per
// pixel instead of one per window (the run-time version above serialises them: 86 us for the ResNet stem pool).
template <typename T, int K, int S>
__global__ void __launch_bounds__(256) maxpool_bwd_fixed_kernel(const T* __restrict__ dy,
                                                                 const uint8_t* __restrict__ code, T* __restrict__ dx,
                                                                 int N, int H, int W, int C, int OH, int OW, int p) {
  constexpr int V = VecIO<T>::N;
  using CV = typename CodeVec<V>::type;
  pdl_wait();
  pdl_trigger();
  const int cv = C / V;
  const size_t total = (size_t)N * H * W * cv;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % cv) * V;
    size_t t = i / cv;
    const int iw = t % W;
    t /= W;
    const int ih = t % H;
    const int n = (int)(t / H);
    float v[K * K][V];
    int cd[K * K][V];
    bool ok[K * K];
#pragma unroll
    for (int kh = 0; kh < K; ++kh) {
      const int th = ih + p - kh;
      const bool okh = th >= 0 && th % S == 0 && th / S < OH;
#pragma unroll
      for (int kw = 0; kw < K; ++kw) {
        const int tw = iw + p - kw;
        const int tap = kh * K + kw;
        ok[tap] = okh && tw >= 0 && tw % S == 0 && tw / S < OW;
        if (ok[tap]) {
          const size_t o = (((size_t)n * OH + th / S) * OW + tw / S) * C + c;
          unpack_code(__ldg(reinterpret_cast<const CV*>(code + o)), cd[tap]);
          VecIO<T>::load(dy + o, v[tap]);
        }
      }
    }
    float g[V];
#pragma unroll
    for (int j = 0; j < V; ++j) g[j] = 0.f;
#pragma unroll
    for (int tap = 0; tap < K * K; ++tap)
      if (ok[tap]) {
#pragma unroll
        for (int j = 0; j < V; ++j)
          if (cd[tap][j] == tap) g[j] += v[tap][j];
      }
    VecIO<T>::store(dx + i * V, g);
  }
}

// Global average pooling: y[n,c] (fp32) = mean over H*W of x; backward: dx[n,h,w,c] = dy[n,c] / (H*W),
// zeroed where msk <= 0 (msk = the pooled tensor itself when it is the output of a ReLU).
template <typename T>
__global__ void __launch_bounds__(256) gap_fwd_kernel(const T* __restrict__ x, float* __restrict__ y, int N, int HW,
                                                       int C) {
  constexpr int V = VecIO<T>::N;
  pdl_wait();
  pdl_trigger();
  const int cv = C / V;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N * cv) return;
  const int n = i / cv, c = (i % cv) * V;
  float s[V];
#pragma unroll
  for (int j = 0; j < V; ++j) s[j] = 0.f;
#pragma unroll 4
  for (int p = 0; p < HW; ++p) {
    float v[V];
    VecIO<T>::load(x + ((size_t)n * HW + p) * C + c, v);
#pragma unroll
    for (int j = 0; j < V; ++j) s[j] += v[j];
  }
  const float inv = 1.f / (float)HW;
#pragma unroll
  for (int j = 0; j < V; ++j) y[(size_t)n * C + c + j] = s[j] * inv;
}

template <typename T>
__global__ void __launch_bounds__(256) gap_bwd_kernel(const float* __restrict__ dy, const T* __restrict__ msk,
                                                       T* __restrict__ dx, int N, int HW, int C) {
  constexpr int V = VecIO<T>::N;
  pdl_wait();
  pdl_trigger();
  const int cv = C / V;
  const size_t total = (size_t)N * HW * cv;
  const float inv = 1.f / (float)HW;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % cv) * V;
    const int n = (int)(i / ((size_t)HW * cv));
    float g[V];
#pragma unroll
    for (int j = 0; j < V; ++j) g[j] = dy[(size_t)n * C + c + j] * inv;
    if (msk != nullptr) {
      float m[V];
      VecIO<T>::load(msk + i * V, m);
#pragma unroll
      for (int j = 0; j < V; ++j) g[j] = m[j] > 0.f ? g[j] : 0.f;
    }
    VecIO<T>::store(dx + i * V, g);
  }
}

// ---------------------------------------------------------------------------------------
// im2col for convolutions whose channel count is too small for the implicit-GEMM tensor-core
// path (a pixel is less than one 128-byte TMA row, e.g. the 3-channel stem of ResNet): the patch
// matrix col[pixel][tap*Ci + ci] (row pitch kpad, padding columns zero) is materialised once per
// step and feeds the tcgen05 GEMM core for the forward product (col . W) and the filter
// gradient (col^T . dy).
// ---------------------------------------------------------------------------------------
struct PatchShape {
  int N, H, W, Ci, KH, KW, OH, OW, sh, sw, ph, pw;
};

template <typename TI, typename TO>
__global__ void __launch_bounds__(256) im2col_kernel(const TI* __restrict__ x, TO* __restrict__ col, PatchShape s,
                                                      int kpad) {
  // One thread = one 16-byte vector of one patch row: VO consecutive k = (kh, kw, ci) entries, walked with
  // carries instead of a division per element; the loads of a vector are independent.
  constexpr int VO = VecIO<TO>::N;
  pdl_wait();
  pdl_trigger();
  const int K = s.KH * s.KW * s.Ci;
  const int kv = kpad / VO;
  const size_t total = (size_t)s.N * s.OH * s.OW * kv;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int k0 = (int)(i % kv) * VO;
    const size_t p = i / kv;
    const int ow = p % s.OW;
    const int oh = (p / s.OW) % s.OH;
    const int n = (int)(p / ((size_t)s.OW * s.OH));
    const int ih0 = oh * s.sh - s.ph, iw0 = ow * s.sw - s.pw;
    int ci = k0 % s.Ci;
    int kw = (k0 / s.Ci) % s.KW;
    int kh = k0 / (s.Ci * s.KW);
    float v[VO];
#pragma unroll
    for (int j = 0; j < VO; ++j) {
      const int ih = ih0 + kh, iw = iw0 + kw;
      v[j] = 0.f;
      if (k0 + j < K && ih >= 0 && ih < s.H && iw >= 0 && iw < s.W)
        v[j] = load_scalar(x + (((size_t)n * s.H + ih) * s.W + iw) * s.Ci + ci);
      if (++ci == s.Ci) {
        ci = 0;
        if (++kw == s.KW) kw = 0, ++kh;
      }
    }
    VecIO<TO>::store(col + p * kpad + k0, v);
  }
}

// dst = (TO) src, 16 bytes of the narrower type per thread and iteration.
template <typename TI, typename TO>
__global__ void __launch_bounds__(256) cast_kernel(const TI* __restrict__ src, TO* __restrict__ dst, size_t n) {
  pdl_wait();
  pdl_trigger();
  const size_t nvec = n / 8;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nvec; i += (size_t)gridDim.x * blockDim.x) {
    float v[8];
    if constexpr (sizeof(TI) == 4) {
      float lo[4], hi[4];
      VecIO<float>::load(reinterpret_cast<const float*>(src) + i * 8, lo);
      VecIO<float>::load(reinterpret_cast<const float*>(src) + i * 8 + 4, hi);
#pragma unroll
      for (int k = 0; k < 4; ++k) v[k] = lo[k], v[4 + k] = hi[k];
    } else {
      VecIO<__nv_bfloat16>::load(reinterpret_cast<const __nv_bfloat16*>(src) + i * 8, v);
    }
    if constexpr (sizeof(TO) == 4) {
      const float lo[4] = {v[0], v[1], v[2], v[3]}, hi[4] = {v[4], v[5], v[6], v[7]};
      VecIO<float>::store(reinterpret_cast<float*>(dst) + i * 8, lo);
      VecIO<float>::store(reinterpret_cast<float*>(dst) + i * 8 + 4, hi);
    } else {
      VecIO<__nv_bfloat16>::store(reinterpret_cast<__nv_bfloat16*>(dst) + i * 8, v);
    }
  }
  // tail (n % 8 elements)
  const size_t t = nvec * 8 + blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (t < n) store_scalar(dst + t, load_scalar(src + t));
}

static inline int grid_1d(size_t n, int block = 256, int ctas_per_sm = 16) {
  const int cap = sm_count() * ctas_per_sm;
  size_t g = (n + block - 1) / block;
  return (int)(g < 1 ? 1 : (g > (size_t)cap ? cap : g));
}

// lanes_c (power of two <= 32 covering C/V channel vectors) and the row split of the per-channel kernels
struct ChanGrid {
  int lanes_c, rows_per_cta;
  dim3 grid;
};
// ctas: how many CTAs to aim for — streaming kernels want many (8 per SM), the reduction kernels few
// and fat ones (2 per SM): every CTA ends in atomics on the same 2C floats.
static inline ChanGrid chan_grid(int M, int C, int V, int ctas = 0) {
  if (ctas <= 0) ctas = sm_count() * 8;
  const int cvec = C / V;
  int lanes = 1;
  while (lanes < 32 && lanes < cvec) lanes <<= 1;
  ChanGrid g;
  g.lanes_c = lanes;
  const int gx = ceil_div(cvec, lanes);
  const int nrg = 256 / lanes;
  const int target_y = max(1, ctas / gx);
  g.rows_per_cta = max(nrg * 4, ceil_div(M, target_y));
  g.grid = dim3(gx, ceil_div(M, g.rows_per_cta));
  return g;
}

}  // namespace b2d

using namespace b2d;
typedef __nv_bfloat16 bf16_t;

#define B2D_DTYPE_OK(dt) B2D_REQUIRE((dt) == B2D_F32 || (dt) == B2D_BF16, "unknown dtype %d", (int)(dt))

static inline int vec_of(int dtype) { return dtype == B2D_F32 ? 4 : 8; }
static inline int reduce_ctas() { return sm_count() * 2; }
// Reduction kernels end in atomics on 2C floats: few fat CTAs when those floats are a handful of L2 lines
// (C <= 128), more CTAs (better latency hiding: the kernels run at ~25 % occupancy) when they spread out.
// B2D_BN_REDUCE_CTAS_PER_SM overrides the per-SM count for wide tensors (default 2 = same as narrow).
static int reduce_ctas(int C) {
  static const int wide = [] {
    const char* e = getenv("B2D_BN_REDUCE_CTAS_PER_SM");
    const int v = e != nullptr ? atoi(e) : 2;
    return v >= 1 && v <= 8 ? v : 2;
  }();
  return C <= 128 ? reduce_ctas() : sm_count() * wide;
}

static int fill_bn_args(const b2d_bn_desc& d, BnArgs* a) {
  B2D_DTYPE_OK(d.dtype);
  const int V = vec_of(d.dtype);
  B2D_REQUIRE(d.M > 0 && d.C > 0 && d.C % V == 0, "batchnorm: C (%d) must be a multiple of %d", d.C, V);
  a->x = d.x, a->y = d.y, a->dy = d.dy, a->dx = d.dx, a->residual = d.residual;
  a->gamma = d.gamma, a->beta = d.beta;
  a->moving_mean = d.moving_mean, a->moving_var = d.moving_var;
  a->moving_mean_grad = d.moving_mean_grad, a->moving_var_grad = d.moving_var_grad;
  a->stats = d.stats, a->saved = d.saved, a->red = d.red, a->dgamma = d.dgamma, a->dbeta = d.dbeta;
  a->M = d.M, a->C = d.C, a->relu = d.relu, a->inference = d.inference;
  a->eps = d.eps, a->momentum = d.momentum, a->inv_lr = d.inv_lr;
  return B2D_OK;
}

extern "C" int b2d_bn_forward(const b2d_bn_desc* dp, void* stream) {
  B2D_REQUIRE(dp != nullptr, "bn_forward: null descriptor");
  const b2d_bn_desc& d = *dp;
  BnArgs a{};
  int rc = fill_bn_args(d, &a);
  if (rc != B2D_OK) return rc;
  B2D_REQUIRE(d.x && d.y && d.gamma && d.beta, "bn_forward: null tensor");
  if (d.inference)
    B2D_REQUIRE(d.moving_mean && d.moving_var, "bn_forward: inference needs the moving statistics");
  else
    B2D_REQUIRE(d.stats && d.saved && d.red, "bn_forward: training needs stats / saved / red scratch");
  B2D_REQUIRE((d.moving_mean_grad == nullptr) == (d.moving_var_grad == nullptr) &&
                  (d.moving_mean_grad == nullptr || (d.moving_mean && d.moving_var)),
              "bn_forward: inconsistent moving-statistics pointers");
  const ChanGrid g = chan_grid(d.M, d.C, vec_of(d.dtype));
  const ChanGrid gr = chan_grid(d.M, d.C, vec_of(d.dtype), reduce_ctas(d.C));
  a.rows_per_cta = g.rows_per_cta, a.lanes_c = g.lanes_c;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (d.dtype == B2D_F32) {
    if (!d.inference)
      B2D_CUDA_CHECK(launch_pdl(col_stats_kernel<float, true>, gr.grid, dim3(256), 0, st, static_cast<const float*>(d.x),
                                d.stats, d.M, d.C, gr.rows_per_cta, gr.lanes_c));
    B2D_CUDA_CHECK(launch_pdl(bn_apply_kernel<float>, g.grid, dim3(256), 0, st, a));
  } else {
    if (!d.inference)
      B2D_CUDA_CHECK(launch_pdl(col_stats_kernel<bf16_t, true>, gr.grid, dim3(256), 0, st,
                                static_cast<const bf16_t*>(d.x), d.stats, d.M, d.C, gr.rows_per_cta, gr.lanes_c));
    B2D_CUDA_CHECK(launch_pdl(bn_apply_kernel<bf16_t>, g.grid, dim3(256), 0, st, a));
  }
  return B2D_OK;
}

extern "C" int b2d_bn_backward(const b2d_bn_desc* dp, void* stream) {
  B2D_REQUIRE(dp != nullptr, "bn_backward: null descriptor");
  const b2d_bn_desc& d = *dp;
  BnArgs a{};
  int rc = fill_bn_args(d, &a);
  if (rc != B2D_OK) return rc;
  B2D_REQUIRE(d.x && d.dy && d.gamma && d.saved && d.red && d.stats && d.dgamma && d.dbeta,
              "bn_backward: null tensor");
  if (d.residual != nullptr) a.relu = 0;   // dy is already the gradient of the pre-activation sum
  B2D_REQUIRE(!a.relu || d.beta, "bn_backward: the ReLU mask is recomputed from x and needs beta");
  const ChanGrid g = chan_grid(d.M, d.C, vec_of(d.dtype));
  const ChanGrid gr = chan_grid(d.M, d.C, vec_of(d.dtype), reduce_ctas(d.C));
  a.rows_per_cta = g.rows_per_cta, a.lanes_c = g.lanes_c;
  BnArgs ar = a;
  ar.rows_per_cta = gr.rows_per_cta;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (d.dtype == B2D_F32) {
    B2D_CUDA_CHECK(launch_pdl(bn_bwd_reduce_kernel<float>, gr.grid, dim3(256), 0, st, ar));
    B2D_CUDA_CHECK(launch_pdl(bn_bwd_apply_kernel<float>, d.dx ? g.grid : dim3(g.grid.x, 1), dim3(256), 0, st, a));
  } else {
    B2D_CUDA_CHECK(launch_pdl(bn_bwd_reduce_kernel<bf16_t>, gr.grid, dim3(256), 0, st, ar));
    B2D_CUDA_CHECK(launch_pdl(bn_bwd_apply_kernel<bf16_t>, d.dx ? g.grid : dim3(g.grid.x, 1), dim3(256), 0, st, a));
  }
  return B2D_OK;
}

extern "C" int b2d_add_act(int dtype, const void* a, const void* b, const void* mask, void* out, size_t n, int mode,
                           void* stream) {
  B2D_DTYPE_OK(dtype);
  const int V = vec_of(dtype);
  B2D_REQUIRE(a && out && mode >= 0 && mode <= 2, "add_act: bad argument");
  B2D_REQUIRE(mode != 2 || mask, "add_act: mode 2 needs a mask tensor");
  B2D_REQUIRE(n % V == 0, "add_act: element count must be a multiple of %d", V);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t nvec = n / V;
  if (dtype == B2D_F32)
    B2D_CUDA_CHECK(launch_pdl(add_act_kernel<float>, dim3(grid_1d(nvec)), dim3(256), 0, st,
                              static_cast<const float*>(a), static_cast<const float*>(b),
                              static_cast<const float*>(mask), static_cast<float*>(out), nvec, mode));
  else
    B2D_CUDA_CHECK(launch_pdl(add_act_kernel<bf16_t>, dim3(grid_1d(nvec)), dim3(256), 0, st,
                              static_cast<const bf16_t*>(a), static_cast<const bf16_t*>(b),
                              static_cast<const bf16_t*>(mask), static_cast<bf16_t*>(out), nvec, mode));
  return B2D_OK;
}

// p = padding at the top / left; OH, OW = output size (<= 0: derive from symmetric padding).
extern "C" int b2d_maxpool_fwd(int dtype, const void* x, void* y, uint8_t* code, int N, int H, int W, int C, int k,
                               int s, int p, int OH, int OW, void* stream) {
  B2D_DTYPE_OK(dtype);
  const int V = vec_of(dtype);
  B2D_REQUIRE(C % V == 0, "maxpool: C (%d) must be a multiple of %d", C, V);
  B2D_REQUIRE(k >= 1 && k <= 15 && s >= 1, "maxpool: bad window / stride");
  if (OH <= 0) OH = (H + 2 * p - k) / s + 1;
  if (OW <= 0) OW = (W + 2 * p - k) / s + 1;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const dim3 grid(grid_1d((size_t)N * OH * OW * (C / V)));
  if (dtype == B2D_F32)
    B2D_CUDA_CHECK(launch_pdl(maxpool_fwd_kernel<float>, grid, dim3(256), 0, st, static_cast<const float*>(x),
                              static_cast<float*>(y), code, N, H, W, C, OH, OW, k, s, p));
  else
    B2D_CUDA_CHECK(launch_pdl(maxpool_fwd_kernel<bf16_t>, grid, dim3(256), 0, st, static_cast<const bf16_t*>(x),
                              static_cast<bf16_t*>(y), code, N, H, W, C, OH, OW, k, s, p));
  return B2D_OK;
}

extern "C" int b2d_maxpool_bwd(int dtype, const void* dy, const uint8_t* code, void* dx, int N, int H, int W, int C,
                               int k, int s, int p, int OH, int OW, void* stream) {
  B2D_DTYPE_OK(dtype);
  const int V = vec_of(dtype);
  B2D_REQUIRE(C % V == 0, "maxpool: C (%d) must be a multiple of %d", C, V);
  if (OH <= 0) OH = (H + 2 * p - k) / s + 1;
  if (OW <= 0) OW = (W + 2 * p - k) / s + 1;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const dim3 grid(grid_1d((size_t)N * H * W * (C / V)));
#define B2D_POOL_BWD_FIXED(T, K, S)                                                                              \
  B2D_CUDA_CHECK(launch_pdl(maxpool_bwd_fixed_kernel<T, K, S>, grid, dim3(256), 0, st, static_cast<const T*>(dy), code, \
                            static_cast<T*>(dx), N, H, W, C, OH, OW, p))
  // B2D_POOL_BWD_FIXED=1 selects the compile-time-window kernels (off until measured on the GPU)
  static const bool fixed = [] {
    const char* e = getenv("B2D_POOL_BWD_FIXED");
    return e != nullptr && e[0] == '1';
  }();
  if (!fixed) k = -k;      // no specialisation matches a negative window
  if (dtype == B2D_F32) {
    if (k == 3 && s == 2) B2D_POOL_BWD_FIXED(float, 3, 2);
    else if (k == 2 && s == 2) B2D_POOL_BWD_FIXED(float, 2, 2);
    else
      B2D_CUDA_CHECK(launch_pdl(maxpool_bwd_kernel<float>, grid, dim3(256), 0, st, static_cast<const float*>(dy), code,
                                static_cast<float*>(dx), N, H, W, C, OH, OW, k < 0 ? -k : k, s, p));
  } else {
    if (k == 3 && s == 2) B2D_POOL_BWD_FIXED(bf16_t, 3, 2);
    else if (k == 2 && s == 2) B2D_POOL_BWD_FIXED(bf16_t, 2, 2);
    else
      B2D_CUDA_CHECK(launch_pdl(maxpool_bwd_kernel<bf16_t>, grid, dim3(256), 0, st, static_cast<const bf16_t*>(dy), code,
                                static_cast<bf16_t*>(dx), N, H, W, C, OH, OW, k < 0 ? -k : k, s, p));
  }
#undef B2D_POOL_BWD_FIXED
  return B2D_OK;
}

extern "C" int b2d_gap_fwd(int dtype, const void* x, float* y, int N, int HW, int C, void* stream) {
  B2D_DTYPE_OK(dtype);
  const int V = vec_of(dtype);
  B2D_REQUIRE(C % V == 0, "global_average_pool: C (%d) must be a multiple of %d", C, V);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const dim3 grid(ceil_div(N * (C / V), 128));
  if (dtype == B2D_F32)
    B2D_CUDA_CHECK(launch_pdl(gap_fwd_kernel<float>, grid, dim3(128), 0, st, static_cast<const float*>(x), y, N, HW, C));
  else
    B2D_CUDA_CHECK(launch_pdl(gap_fwd_kernel<bf16_t>, grid, dim3(128), 0, st, static_cast<const bf16_t*>(x), y, N, HW, C));
  return B2D_OK;
}

extern "C" int b2d_gap_bwd(int dtype, const float* dy, const void* mask, void* dx, int N, int HW, int C,
                           void* stream) {
  B2D_DTYPE_OK(dtype);
  const int V = vec_of(dtype);
  B2D_REQUIRE(C % V == 0, "global_average_pool: C (%d) must be a multiple of %d", C, V);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const dim3 grid(grid_1d((size_t)N * HW * (C / V)));
  if (dtype == B2D_F32)
    B2D_CUDA_CHECK(launch_pdl(gap_bwd_kernel<float>, grid, dim3(256), 0, st, dy, static_cast<const float*>(mask),
                              static_cast<float*>(dx), N, HW, C));
  else
    B2D_CUDA_CHECK(launch_pdl(gap_bwd_kernel<bf16_t>, grid, dim3(256), 0, st, dy, static_cast<const bf16_t*>(mask),
                              static_cast<bf16_t*>(dx), N, HW, C));
  return B2D_OK;
}

extern "C" int b2d_im2col(int in_dtype, int out_dtype, const void* x, void* col, int N, int H, int W, int Ci, int KH,
                          int KW, int sh, int sw, int ph, int pw, int OH, int OW, int kpad, void* stream) {
  B2D_DTYPE_OK(in_dtype);
  B2D_DTYPE_OK(out_dtype);
  B2D_REQUIRE(!(in_dtype == B2D_BF16 && out_dtype == B2D_F32), "im2col: bf16 -> fp32 is not provided");
  PatchShape s;
  s.N = N, s.H = H, s.W = W, s.Ci = Ci, s.KH = KH, s.KW = KW, s.sh = sh, s.sw = sw, s.ph = ph, s.pw = pw;
  s.OH = OH > 0 ? OH : (H + 2 * ph - KH) / sh + 1;
  s.OW = OW > 0 ? OW : (W + 2 * pw - KW) / sw + 1;
  B2D_REQUIRE(kpad >= KH * KW * Ci && kpad % vec_of(out_dtype) == 0,
              "im2col: kpad must be >= KH*KW*Ci and keep rows 16-byte aligned");
  const size_t npix = (size_t)N * s.OH * s.OW;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const dim3 grid(grid_1d(npix * (kpad / vec_of(out_dtype))));
  if (in_dtype == B2D_F32 && out_dtype == B2D_F32)
    B2D_CUDA_CHECK(launch_pdl(im2col_kernel<float, float>, grid, dim3(256), 0, st, static_cast<const float*>(x),
                              static_cast<float*>(col), s, kpad));
  else if (in_dtype == B2D_F32)
    B2D_CUDA_CHECK(launch_pdl(im2col_kernel<float, bf16_t>, grid, dim3(256), 0, st, static_cast<const float*>(x),
                              static_cast<bf16_t*>(col), s, kpad));
  else
    B2D_CUDA_CHECK(launch_pdl(im2col_kernel<bf16_t, bf16_t>, grid, dim3(256), 0, st, static_cast<const bf16_t*>(x),
                              static_cast<bf16_t*>(col), s, kpad));
  return B2D_OK;
}

extern "C" int b2d_cast(int src_dtype, int dst_dtype, const void* src, void* dst, size_t n, void* stream) {
  B2D_DTYPE_OK(src_dtype);
  B2D_DTYPE_OK(dst_dtype);
  B2D_REQUIRE(src_dtype != dst_dtype, "cast: source and destination types are the same");
  B2D_REQUIRE((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) % 16 == 0,
              "cast: pointers must be 16-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const dim3 grid(grid_1d(n / 8 + 8));
  if (src_dtype == B2D_F32)
    B2D_CUDA_CHECK(launch_pdl(cast_kernel<float, bf16_t>, grid, dim3(256), 0, st, static_cast<const float*>(src),
                              static_cast<bf16_t*>(dst), n));
  else
    B2D_CUDA_CHECK(launch_pdl(cast_kernel<bf16_t, float>, grid, dim3(256), 0, st, static_cast<const bf16_t*>(src),
                              static_cast<float*>(dst), n));
  return B2D_OK;
}

// db[0:C] += column sums of dy [M, C] (bias gradient), any supported storage type.
extern "C" int b2d_colsum_t(int dtype, const void* dy, float* db, int M, int C, void* stream) {
  B2D_DTYPE_OK(dtype);
  const int V = vec_of(dtype);
  B2D_REQUIRE(C % V == 0, "colsum: C (%d) must be a multiple of %d", C, V);
  const ChanGrid g = chan_grid(M, C, V, reduce_ctas(C));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == B2D_F32)
    B2D_CUDA_CHECK(launch_pdl(col_stats_kernel<float, false>, g.grid, dim3(256), 0, st, static_cast<const float*>(dy), db,
                              M, C, g.rows_per_cta, g.lanes_c));
  else
    B2D_CUDA_CHECK(launch_pdl(col_stats_kernel<bf16_t, false>, g.grid, dim3(256), 0, st,
                              static_cast<const bf16_t*>(dy), db, M, C, g.rows_per_cta, g.lanes_c));
  return B2D_OK;
}
