// nn_kernels.cu — non-GEMM kernels of the training step (sm_100a, fp32 storage).
//
//   * first-layer convolution (C_in = 1) fused with bias + ReLU + 2x2 max-pool, forward and
//     the matching fused pool-backward + ReLU-backward + filter/bias gradient
//   * classifier head: Dense(C) + softmax cross-entropy + all of their gradients in one kernel
//   * stand-alone softmax cross-entropy (+gradient, +accuracy count)
//   * 2x2 max-pool forward / backward, bias gradient (column sums), ReLU backward
//   * direct (CUDA-core) convolution fallbacks for shapes the tensor-core path does not take
//   * SGD apply (+gradient zeroing) for the single-replica case
//
// Layouts follow Keras: activations NHWC, conv kernels HWIO, Dense kernels (in, out).
// Reference behaviour: README.md:292-302 (the layer list, loss, optimizer and metric).
#include "host_common.h"
#include "ptx.cuh"

namespace b2d {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// =======================================================================================
// Conv2D(C_in = 1, KxK, valid, stride 1) + bias + ReLU + MaxPool(2x2)        [K1 + K2]
//   one CTA = one image x a band of pooled rows; lane = output channel; the image band sits
//   in shared memory and is read by broadcast; the 26x26xCo conv map never reaches HBM.
//   code[n,ph,pw,c] = argmax position (0..3 = dy*2+dx) or 4 when ReLU clamped the maximum.
// =======================================================================================
//   CG = channel groups of 32 handled per pixel: the (KS+1)^2 input window is fetched once (broadcast
//   shared-memory loads are the bottleneck of this kernel) and reused for CG x 32 output channels.
template <int KS, int CG>
__global__ void __launch_bounds__(256) conv_c1_pool_fwd_kernel(
    const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
    float* __restrict__ y, uint8_t* __restrict__ code, int H, int W, int Co, int PH, int PW,
    int band) {
  extern __shared__ float xs[];  // [(2*band + KS) rows][W]
  const int n = blockIdx.x;
  const int ph0 = blockIdx.y * band;
  const int ph1 = min(PH, ph0 + band);
  const int r0 = 2 * ph0;
  const int r1 = min(H, 2 * ph1 + KS - 1);
  const float* xi = x + (size_t)n * H * W;
  // The input batch was copied in before this step's graph started (an earlier operation than the kernel this
  // one is launched behind), so staging it does not have to wait for the predecessor: with programmatic
  // dependent launch this overlaps the tail of the previous step's update.  The weights do.
  for (int i = threadIdx.x; i < (r1 - r0) * W; i += blockDim.x) xs[i] = xi[r0 * W + i];
  pdl_wait();
  pdl_trigger();
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarp = blockDim.x >> 5;
  const int npix = (ph1 - ph0) * PW;
  for (int cg0 = 0; cg0 < Co; cg0 += 32 * CG) {
    float wr[CG][KS * KS], bv[CG];
#pragma unroll
    for (int g = 0; g < CG; ++g) {
      const int c = cg0 + 32 * g + lane;
#pragma unroll
      for (int t = 0; t < KS * KS; ++t) wr[g][t] = c < Co ? w[t * Co + c] : 0.f;
      bv[g] = c < Co ? bias[c] : 0.f;
    }
    for (int p = warp; p < npix; p += nwarp) {
      const int ph = ph0 + p / PW, pw = p % PW;
      const float* base = xs + (2 * ph - r0) * W + 2 * pw;
      float win[(KS + 1) * (KS + 1)];
#pragma unroll
      for (int i = 0; i < KS + 1; ++i)
#pragma unroll
        for (int j = 0; j < KS + 1; ++j) win[i * (KS + 1) + j] = base[i * W + j];
#pragma unroll
      for (int g = 0; g < CG; ++g) {
        const int c = cg0 + 32 * g + lane;
        float best = -INFINITY;
        int arg = 0;
#pragma unroll
        for (int dy = 0; dy < 2; ++dy)
#pragma unroll
          for (int dx = 0; dx < 2; ++dx) {
            float acc = bv[g];
#pragma unroll
            for (int i = 0; i < KS; ++i)
#pragma unroll
              for (int j = 0; j < KS; ++j) acc = fmaf(win[(dy + i) * (KS + 1) + dx + j], wr[g][i * KS + j], acc);
            if (acc > best) best = acc, arg = dy * 2 + dx;
          }
        if (c < Co) {
          const size_t o = (((size_t)n * PH + ph) * PW + pw) * Co + c;
          y[o] = fmaxf(best, 0.f);
          code[o] = best > 0.f ? (uint8_t)arg : (uint8_t)4;
        }
      }
    }
  }
}

// Fused MaxPool-backward + ReLU-backward + conv filter/bias gradient for the layer above
// [K8 + K9]: dW[i,j,0,c] += sum g * x[n, 2ph+ay+i, 2pw+ax+j],  db[c] += sum g.
template <int KS>
__global__ void __launch_bounds__(256) conv_c1_pool_bwd_kernel(
    const float* __restrict__ x, const float* __restrict__ dy, const uint8_t* __restrict__ code,
    float* __restrict__ dw, float* __restrict__ db, int H, int W, int Co, int PH, int PW, int band) {
  extern __shared__ float sm[];
  float* xs = sm;  // image band
  const int n = blockIdx.x;
  const int ph0 = blockIdx.y * band;
  const int ph1 = min(PH, ph0 + band);
  const int r0 = 2 * ph0;
  const int r1 = min(H, 2 * ph1 + KS - 1);
  const int band_elems = (2 * band + KS) * W;
  float* red = sm + band_elems;  // [nwarp][KS*KS+1][32]
  const float* xi = x + (size_t)n * H * W;
  // the input images do not depend on the predecessor kernel (see conv_c1_pool_fwd_kernel): stage them first
  for (int i = threadIdx.x; i < (r1 - r0) * W; i += blockDim.x) xs[i] = xi[r0 * W + i];
  pdl_wait();
  pdl_trigger();
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarp = blockDim.x >> 5;
  const int npix = (ph1 - ph0) * PW;
  constexpr int NT = KS * KS + 1;
  for (int cg = 0; cg < Co; cg += 32) {
    const int c = cg + lane;
    const bool c_ok = c < Co;
    float acc[NT];
#pragma unroll
    for (int t = 0; t < NT; ++t) acc[t] = 0.f;
    // Eight pooled pixels per warp and round: their (gradient, arg-max code) pairs are fetched together —
    // a warp's whole share of a band in one or two memory round trips instead of one per pixel.
    constexpr int PF = 8;
    for (int p0 = warp; p0 < npix; p0 += nwarp * PF) {
      float g[PF];
      int cd[PF];
#pragma unroll
      for (int u = 0; u < PF; ++u) {
        const int p = p0 + u * nwarp;
        const bool ok = c_ok && p < npix;
        const int pc = ok ? p : 0;
        const size_t o = (((size_t)n * PH + ph0 + pc / PW) * PW + pc % PW) * Co + (c_ok ? c : 0);
        g[u] = ok ? __ldg(dy + o) : 0.f;
        cd[u] = ok ? (int)__ldg(code + o) : 4;
      }
#pragma unroll
      for (int u = 0; u < PF; ++u) {
        if (cd[u] < 4) {
          const int p = p0 + u * nwarp;
          const int ph = ph0 + p / PW, pw = p % PW;
          const float* base = xs + (2 * ph - r0 + (cd[u] >> 1)) * W + 2 * pw + (cd[u] & 1);
#pragma unroll
          for (int i = 0; i < KS; ++i)
#pragma unroll
            for (int j = 0; j < KS; ++j) acc[i * KS + j] = fmaf(g[u], base[i * W + j], acc[i * KS + j]);
          acc[KS * KS] += g[u];
        }
      }
    }
#pragma unroll
    for (int t = 0; t < NT; ++t) red[(warp * NT + t) * 32 + lane] = acc[t];
    __syncthreads();
    for (int i = threadIdx.x; i < NT * 32; i += blockDim.x) {
      const int t = i / 32, l = i % 32;
      float s = 0.f;
      for (int wv = 0; wv < nwarp; ++wv) s += red[(wv * NT + t) * 32 + l];
      const int cc = cg + l;
      if (cc < Co) {
        if (t < KS * KS)
          atomicAdd(&dw[t * Co + cc], s);
        else
          atomicAdd(&db[cc], s);
      }
    }
    __syncthreads();
  }
}

// =======================================================================================
// Classifier head: logits = h.W + b ; softmax cross-entropy ; all gradients.   [K4+K5+K6]
//   Phase 1 — one warp per batch row (8 rows per pass): class c's logit is a warp-reduced dot
//   product against W in shared memory; the softmax runs with lane = class; dh = dz.W^T
//   (ReLU-masked) is written straight out.
//   Phase 2 — the whole CTA turns the pass's 8 (h, dz) rows into dW / db partial sums; each
//   thread owns fixed (k, c) outputs, so there are no shared-memory atomics; one global atomic
//   per output per CTA.  Optionally also accumulates the bias gradient of the layer that
//   produced h (column sums of dh), saving that layer a kernel.
//   The kernel runs once per step on a handful of CTAs, i.e. it is instruction-fetch bound:
//   every loop is deliberately NOT unrolled to keep the code a few hundred instructions.
//   metrics[0] += sum of per-row losses, metrics[1] += #correct, metrics[2] += #rows.
// =======================================================================================
constexpr int kHeadMaxC = 32;
constexpr int kHeadRows = 8;  // rows per pass = warps per CTA

__global__ void __launch_bounds__(256) dense_xent_head_kernel(
    const float* __restrict__ h, const float* __restrict__ w, const float* __restrict__ bias,
    const int* __restrict__ labels, float* __restrict__ logits_out, float* __restrict__ dh,
    float* __restrict__ dw, float* __restrict__ db, float* __restrict__ dh_colsum,
    float* __restrict__ metrics, int B, int K, int C, float grad_scale, int relu_mask,
    int rows_per_cta) {
  extern __shared__ float sm[];
  const int KC = K * C;
  float* ws = sm;                             // [K][C]
  float* hs = ws + KC;                        // [8][K]   h rows of this pass
  float* dhs = hs + kHeadRows * K;            // [8][K]   dh rows of this pass
  float* dzs = dhs + kHeadRows * K;           // [8][32]  dz rows of this pass
  float* rowm = dzs + kHeadRows * kHeadMaxC;  // [8][2]   loss, correct
  float* accw = rowm + 2 * kHeadRows;         // [K*C]    multi-pass dW partial sums
  float* acck = accw + KC;                    // [K]      multi-pass dh column sums
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int row0 = blockIdx.x * rows_per_cta;
  const int row1 = min(B, row0 + rows_per_cta);
  const bool multi_pass = row1 - row0 > kHeadRows;
  // This layer's weights were last written by the previous step's update, which completed before the
  // predecessor of this launch could even start: staging them overlaps the predecessor's tail.
#pragma unroll 1
  for (int i = tid; i < KC; i += 256) ws[i] = w[i], accw[i] = 0.f;
#pragma unroll 1
  for (int i = tid; i < K; i += 256) acck[i] = 0.f;
  float acc_b = 0.f, loss_sum = 0.f, correct_sum = 0.f;
  pdl_wait();
  pdl_trigger();
  __syncthreads();
#pragma unroll 1
  for (int base = row0; base < row1; base += kHeadRows) {
    const int r = base + warp;
    float* hrow = hs + warp * K;
    float* dhrow = dhs + warp * K;
    float* dzrow = dzs + warp * kHeadMaxC;
    if (r < row1) {
      const float* hr = h + (size_t)r * K;
#pragma unroll 1
      for (int k = lane; k < K; k += 32) hrow[k] = hr[k];
      __syncwarp();
      // logits: lane c ends up holding z[c]
      float zc = -INFINITY;
#pragma unroll 1
      for (int c = 0; c < C; ++c) {
        float p = 0.f;
#pragma unroll 1
        for (int k = lane; k < K; k += 32) p = fmaf(hrow[k], ws[k * C + c], p);
        p = warp_sum(p);
        if (lane == c) zc = p + bias[c];
      }
      if (logits_out != nullptr && lane < C) logits_out[(size_t)r * C + lane] = zc;
      // softmax over lanes, arg-max with lowest-index tie-break
      float mx = zc;
      int arg = lane < C ? lane : 0x7fffffff;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float om = __shfl_xor_sync(0xffffffffu, mx, o);
        const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
        if (om > mx || (om == mx && oa < arg)) mx = om, arg = oa;
      }
      const float e = lane < C ? __expf(zc - mx) : 0.f;
      const float se = warp_sum(e);
      const int yl = labels[r];
      const float zy = __shfl_sync(0xffffffffu, zc, yl);
      const float dz = lane < C ? (e / se - (lane == yl ? 1.f : 0.f)) * grad_scale : 0.f;
      dzrow[lane] = dz;
      if (lane == 0) rowm[warp * 2] = mx + __logf(se) - zy, rowm[warp * 2 + 1] = arg == yl ? 1.f : 0.f;
      __syncwarp();
#pragma unroll 1
      for (int k = lane; k < K; k += 32) {
        float g = 0.f;
#pragma unroll 1
        for (int c = 0; c < C; ++c) g = fmaf(dzrow[c], ws[k * C + c], g);
        if (relu_mask && !(hrow[k] > 0.f)) g = 0.f;
        dhrow[k] = g;
        if (dh != nullptr) dh[(size_t)r * K + k] = g;
      }
    } else {
#pragma unroll 1
      for (int k = lane; k < K; k += 32) hrow[k] = 0.f, dhrow[k] = 0.f;
      dzrow[lane] = 0.f;
      if (lane == 0) rowm[warp * 2] = 0.f, rowm[warp * 2 + 1] = 0.f;
    }
    __syncthreads();
    // ---- phase 2: CTA-wide partial sums over the 8 rows of this pass ----
#pragma unroll 1
    for (int o = tid; o < KC; o += 256) {
      const int k = o / C, c = o - k * C;
      float s = 0.f;
#pragma unroll
      for (int rr = 0; rr < kHeadRows; ++rr) s = fmaf(hs[rr * K + k], dzs[rr * kHeadMaxC + c], s);
      if (multi_pass)
        accw[o] += s;
      else
        atomicAdd(&dw[o], s);
    }
    if (dh_colsum != nullptr) {
#pragma unroll 1
      for (int k = tid; k < K; k += 256) {
        float s = 0.f;
#pragma unroll
        for (int rr = 0; rr < kHeadRows; ++rr) s += dhs[rr * K + k];
        if (multi_pass)
          acck[k] += s;
        else
          atomicAdd(&dh_colsum[k], s);
      }
    }
    if (tid < C) {
#pragma unroll
      for (int rr = 0; rr < kHeadRows; ++rr) acc_b += dzs[rr * kHeadMaxC + tid];
    }
    if (tid == 0) {
#pragma unroll
      for (int rr = 0; rr < kHeadRows; ++rr) loss_sum += rowm[rr * 2], correct_sum += rowm[rr * 2 + 1];
    }
    __syncthreads();
  }
  if (multi_pass) {
#pragma unroll 1
    for (int o = tid; o < KC; o += 256) atomicAdd(&dw[o], accw[o]);
    if (dh_colsum != nullptr) {
#pragma unroll 1
      for (int k = tid; k < K; k += 256) atomicAdd(&dh_colsum[k], acck[k]);
    }
  }
  if (tid < C) atomicAdd(&db[tid], acc_b);
  if (tid == 0 && metrics != nullptr) {
    atomicAdd(&metrics[0], loss_sum);
    atomicAdd(&metrics[1], correct_sum);
    atomicAdd(&metrics[2], (float)(row1 - row0));
  }
}


// The same head for LARGE batches (scaled configurations: batch 1024 per GPU, K = 1024): throughput- instead
// of instruction-fetch-bound.  One warp per row, 8 rows per pass.  Differences to the kernel above: the class
// loop is innermost and fully unrolled (kC = 16 accumulators in registers: 16 independent FMA chains per k),
// W sits in shared memory with a 16-byte aligned row pitch and is read as float4 (4 shared loads per 16 FMAs
// instead of 1 per FMA), and in the dW phase each thread owns whole rows of dW (all kC classes of its k's), so
// one h value and four broadcast float4 of dz feed 16 FMAs; the partial sums stay in registers across passes
// (one global atomic per output per CTA).
template <int kC>
__global__ void __launch_bounds__(256) dense_xent_head_wide_kernel(
    const float* __restrict__ h, const float* __restrict__ w, const float* __restrict__ bias,
    const int* __restrict__ labels, float* __restrict__ logits_out, float* __restrict__ dh,
    float* __restrict__ dw, float* __restrict__ db, float* __restrict__ dh_colsum,
    float* __restrict__ metrics, int B, int K, int C, float grad_scale, int relu_mask, int passes) {
  static_assert(kC == 16, "the float4 paths assume 16 classes per row");
  constexpr int kRows = 8, kPitch = kC + 4, kOwnK = 4;       // K <= 256 * kOwnK
  extern __shared__ __align__(16) float sm[];
  float* ws = sm;                          // [K][kPitch]
  float* hs = ws + (size_t)K * kPitch;     // [8][K]
  float* dhs = hs + kRows * K;             // [8][K]
  float* dzs = dhs + kRows * K;            // [8][kC]
  float* rowm = dzs + kRows * kC;          // [8][2]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
#pragma unroll 8
  for (int i = tid; i < K * kC; i += 256) {
    const int k = i >> 4, c = i & 15;
    ws[k * kPitch + c] = c < C ? __ldg(w + (size_t)k * C + c) : 0.f;
  }
  float accw[kOwnK][kC];                   // dW rows k = tid + 256 j
  float acck[kOwnK];                       // dh column sums of the same k's
#pragma unroll
  for (int j = 0; j < kOwnK; ++j) {
    acck[j] = 0.f;
#pragma unroll
    for (int c = 0; c < kC; ++c) accw[j][c] = 0.f;
  }
  float acc_b = 0.f, loss_sum = 0.f, correct_sum = 0.f;
  // W was staged above without waiting: it was written by the previous step's update, which completed before the
  // kernel this one is launched behind could start (programmatic dependent launch)
  pdl_wait();
  pdl_trigger();
  __syncthreads();
  const int row0 = blockIdx.x * (kRows * passes);
#pragma unroll 1
  for (int pass = 0; pass < passes; ++pass) {
    const int r = row0 + pass * kRows + warp;
    float* hrow = hs + warp * K;
    float* dhrow = dhs + warp * K;
    float* dzrow = dzs + warp * kC;
    if (r < B) {
      const float* hr = h + (size_t)r * K;
      float p[kC];
#pragma unroll
      for (int c = 0; c < kC; ++c) p[c] = 0.f;
      // the whole row first (K <= 1024: 32 loads in flight per lane), then the arithmetic: one memory latency
      // per row instead of one per k
      float hreg[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) hreg[j] = lane + 32 * j < K ? __ldg(hr + lane + 32 * j) : 0.f;
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const int k = lane + 32 * j;
        if (k >= K) continue;
        const float hv = hreg[j];
        hrow[k] = hv;
        const float4* wr = reinterpret_cast<const float4*>(ws + k * kPitch);
#pragma unroll
        for (int q = 0; q < kC / 4; ++q) {
          const float4 w4 = wr[q];
          p[4 * q + 0] = fmaf(hv, w4.x, p[4 * q + 0]), p[4 * q + 1] = fmaf(hv, w4.y, p[4 * q + 1]);
          p[4 * q + 2] = fmaf(hv, w4.z, p[4 * q + 2]), p[4 * q + 3] = fmaf(hv, w4.w, p[4 * q + 3]);
        }
      }
      float zc = -INFINITY;
#pragma unroll
      for (int c = 0; c < kC; ++c) {
        const float t = warp_sum(p[c]);
        if (lane == c && c < C) zc = t + bias[c];
      }
      if (logits_out != nullptr && lane < C) logits_out[(size_t)r * C + lane] = zc;
      float mx = zc;
      int arg = lane < C ? lane : 0x7fffffff;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float om = __shfl_xor_sync(0xffffffffu, mx, o);
        const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
        if (om > mx || (om == mx && oa < arg)) mx = om, arg = oa;
      }
      const float e = lane < C ? __expf(zc - mx) : 0.f;
      const float se = warp_sum(e);
      const int yl = labels[r];
      const float zy = __shfl_sync(0xffffffffu, zc, yl);
      const float dz = lane < C ? (e / se - (lane == yl ? 1.f : 0.f)) * grad_scale : 0.f;
      if (lane < kC) dzrow[lane] = dz;
      if (lane == 0) rowm[warp * 2] = mx + __logf(se) - zy, rowm[warp * 2 + 1] = arg == yl ? 1.f : 0.f;
      float dzr[kC];
#pragma unroll
      for (int c = 0; c < kC; ++c) dzr[c] = __shfl_sync(0xffffffffu, dz, c);
#pragma unroll 2
      for (int k = lane; k < K; k += 32) {
        const float4* wr = reinterpret_cast<const float4*>(ws + k * kPitch);
        float g = 0.f;
#pragma unroll
        for (int q = 0; q < kC / 4; ++q) {
          const float4 w4 = wr[q];
          g = fmaf(dzr[4 * q + 0], w4.x, g), g = fmaf(dzr[4 * q + 1], w4.y, g);
          g = fmaf(dzr[4 * q + 2], w4.z, g), g = fmaf(dzr[4 * q + 3], w4.w, g);
        }
        if (relu_mask && !(hrow[k] > 0.f)) g = 0.f;
        dhrow[k] = g;
        if (dh != nullptr) dh[(size_t)r * K + k] = g;
      }
    } else {
      for (int k = lane; k < K; k += 32) hrow[k] = 0.f, dhrow[k] = 0.f;
      if (lane < kC) dzrow[lane] = 0.f;
      if (lane == 0) rowm[warp * 2] = 0.f, rowm[warp * 2 + 1] = 0.f;
    }
    __syncthreads();
    // ---- dW / column sums: thread owns rows k = tid + 256 j of dW ----
#pragma unroll
    for (int rr = 0; rr < kRows; ++rr) {
      float dzv[kC];
#pragma unroll
      for (int q = 0; q < kC / 4; ++q) {
        const float4 t4 = reinterpret_cast<const float4*>(dzs + rr * kC)[q];     // broadcast
        dzv[4 * q] = t4.x, dzv[4 * q + 1] = t4.y, dzv[4 * q + 2] = t4.z, dzv[4 * q + 3] = t4.w;
      }
#pragma unroll
      for (int j = 0; j < kOwnK; ++j) {
        const int k = tid + 256 * j;
        if (k < K) {
          const float hv = hs[rr * K + k];
#pragma unroll
          for (int c = 0; c < kC; ++c) accw[j][c] = fmaf(hv, dzv[c], accw[j][c]);
          acck[j] += dhs[rr * K + k];
        }
      }
    }
    if (tid < C) {
#pragma unroll
      for (int rr = 0; rr < kRows; ++rr) acc_b += dzs[rr * kC + tid];
    }
    if (tid == 0) {
#pragma unroll
      for (int rr = 0; rr < kRows; ++rr) loss_sum += rowm[rr * 2], correct_sum += rowm[rr * 2 + 1];
    }
    __syncthreads();
  }
#pragma unroll
  for (int j = 0; j < kOwnK; ++j) {
    const int k = tid + 256 * j;
    if (k < K) {
      float* dst = dw + (size_t)k * C;
      if ((C & 1) == 0) {          // rows of dW start 8-byte aligned: half as many reductions, as float2
#pragma unroll
        for (int c = 0; c < kC; c += 2)
          if (c < C)
            asm volatile("red.relaxed.gpu.global.add.v2.f32 [%0], {%1, %2};" ::"l"(dst + c), "f"(accw[j][c]),
                         "f"(accw[j][c + 1])
                         : "memory");
      } else {
#pragma unroll
        for (int c = 0; c < kC; ++c)
          if (c < C) atomicAdd(dst + c, accw[j][c]);
      }
      if (dh_colsum != nullptr) atomicAdd(&dh_colsum[k], acck[j]);
    }
  }
  if (tid < C) atomicAdd(&db[tid], acc_b);
  if (tid == 0 && metrics != nullptr) {
    atomicAdd(&metrics[0], loss_sum);
    atomicAdd(&metrics[1], correct_sum);
    atomicAdd(&metrics[2], (float)(min(B, row0 + kRows * passes) - row0));
  }
}

// Stand-alone softmax cross-entropy on logits [B][C] (any C): one warp per row.   [K5]
__global__ void __launch_bounds__(256) softmax_xent_kernel(const float* __restrict__ z,
                                                            const int* __restrict__ labels,
                                                            float* __restrict__ dz,
                                                            float* __restrict__ metrics, int B, int C,
                                                            float grad_scale) {
  pdl_wait();
  pdl_trigger();
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= B) return;
  const float* zr = z + (size_t)warp * C;
  float mx = -INFINITY;
  int arg = 0x7fffffff;
  for (int c = lane; c < C; c += 32) {
    const float v = zr[c];
    if (v > mx) mx = v, arg = c;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {  // argmax with lowest-index tie-break
    const float om = __shfl_xor_sync(0xffffffffu, mx, o);
    const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
    if (om > mx || (om == mx && oa < arg)) mx = om, arg = oa;
  }
  float se = 0.f;
  for (int c = lane; c < C; c += 32) se += __expf(zr[c] - mx);
  se = warp_sum(se);
  const int yl = labels[warp];
  const float inv = 1.f / se;
  if (dz != nullptr)
    for (int c = lane; c < C; c += 32)
      dz[(size_t)warp * C + c] = (__expf(zr[c] - mx) * inv - (c == yl ? 1.f : 0.f)) * grad_scale;
  if (lane == 0 && metrics != nullptr) {
    atomicAdd(&metrics[0], mx + __logf(se) - zr[yl]);
    atomicAdd(&metrics[1], arg == yl ? 1.f : 0.f);
    atomicAdd(&metrics[2], 1.f);
  }
}

// =======================================================================================
// MaxPool 2x2 / stride 2 (NHWC), forward with argmax code and backward (optionally fused with
// the ReLU backward of the producing layer: act_relu => gradient is dropped where y <= 0).
// =======================================================================================
__global__ void maxpool2_fwd_kernel(const float* __restrict__ x, float* __restrict__ y,
                                    uint8_t* __restrict__ code, int N, int H, int W, int C, int PH,
                                    int PW) {
  pdl_wait();
  pdl_trigger();
  const size_t total = (size_t)N * PH * PW * C;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x) {
    const int c = i % C;
    size_t t = i / C;
    const int pw = t % PW;
    t /= PW;
    const int ph = t % PH;
    const int n = t / PH;
    const float* b = x + (((size_t)n * H + 2 * ph) * W + 2 * pw) * C + c;
    float best = b[0];
    int arg = 0;
    float v = b[C];
    if (v > best) best = v, arg = 1;
    v = b[(size_t)W * C];
    if (v > best) best = v, arg = 2;
    v = b[(size_t)W * C + C];
    if (v > best) best = v, arg = 3;
    y[i] = best;
    code[i] = (uint8_t)arg;
  }
}

__global__ void maxpool2_bwd_kernel(const float* __restrict__ dy, const uint8_t* __restrict__ code,
                                    const float* __restrict__ y_pooled, float* __restrict__ dx, int N,
                                    int H, int W, int C, int PH, int PW, int act_relu) {
  pdl_wait();
  pdl_trigger();
  const size_t total = (size_t)N * H * W * C;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x) {
    const int c = i % C;
    size_t t = i / C;
    const int w = t % W;
    t /= W;
    const int hh = t % H;
    const int n = t / H;
    const int ph = hh >> 1, pw = w >> 1;
    float g = 0.f;
    if (ph < PH && pw < PW) {
      const size_t o = (((size_t)n * PH + ph) * PW + pw) * C + c;
      if (code[o] == ((hh & 1) * 2 + (w & 1)) && !(act_relu && !(y_pooled[o] > 0.f))) g = dy[o];
    }
    dx[i] = g;
  }
}

// Same, four channels per thread (C % 4 == 0): 16-byte loads/stores, one index decode per 4 outputs.
__global__ void __launch_bounds__(256) maxpool2_bwd_v4_kernel(const float4* __restrict__ dy,
                                                               const uchar4* __restrict__ code,
                                                               const float4* __restrict__ y_pooled,
                                                               float4* __restrict__ dx, int N, int H, int W,
                                                               int C4, int PH, int PW, int act_relu) {
  pdl_wait();
  pdl_trigger();
  const size_t total = (size_t)N * H * W * C4;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x) {
    const int c = i % C4;
    size_t t = i / C4;
    const int w = t % W;
    t /= W;
    const int hh = t % H;
    const int n = t / H;
    const int ph = hh >> 1, pw = w >> 1;
    float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
    if (ph < PH && pw < PW) {
      const size_t o = (((size_t)n * PH + ph) * PW + pw) * C4 + c;
      const uchar4 cd = code[o];
      const float4 d = dy[o];
      const unsigned char me = (unsigned char)((hh & 1) * 2 + (w & 1));
      float4 yv = make_float4(1.f, 1.f, 1.f, 1.f);
      if (act_relu) yv = y_pooled[o];
      g.x = (cd.x == me && yv.x > 0.f) ? d.x : 0.f;
      g.y = (cd.y == me && yv.y > 0.f) ? d.y : 0.f;
      g.z = (cd.z == me && yv.z > 0.f) ? d.z : 0.f;
      g.w = (cd.w == me && yv.w > 0.f) ? d.w : 0.f;
    }
    dx[i] = g;
  }
}

// Column sums: db[n] += sum_m dy[m][n]   (bias gradient)                         [K6/K7]
__global__ void __launch_bounds__(256) colsum_kernel(const float* __restrict__ dy,
                                                      float* __restrict__ db, int M, int N,
                                                      int rows_per_cta) {
  pdl_wait();
  pdl_trigger();
  __shared__ float part[8][33];
  const int col = blockIdx.x * 32 + (threadIdx.x & 31);
  const int wy = threadIdx.x >> 5;
  const int r0 = blockIdx.y * rows_per_cta, r1 = min(M, r0 + rows_per_cta);
  float s = 0.f;
  if (col < N)
    for (int r = r0 + wy; r < r1; r += 8) s += dy[(size_t)r * N + col];
  part[wy][threadIdx.x & 31] = s;
  __syncthreads();
  if (wy == 0 && col < N) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += part[k][threadIdx.x];
    atomicAdd(&db[col], t);
  }
}

// Vectorised column sums for N % 4 == 0: a warp reads one 512-byte row segment (32 x float4)
// per step, 8 warps walk different rows; partial sums meet in shared memory, one atomic per column.
__global__ void __launch_bounds__(256) colsum_v4_kernel(const float* __restrict__ dy, float* __restrict__ db,
                                                         int M, int N, int rows_per_cta) {
  pdl_wait();
  pdl_trigger();
  __shared__ float4 part[8][32];
  const int lane = threadIdx.x & 31, wy = threadIdx.x >> 5;
  const int col = blockIdx.x * 128 + lane * 4;
  const int r0 = blockIdx.y * rows_per_cta, r1 = min(M, r0 + rows_per_cta);
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  if (col < N)
    for (int r = r0 + wy; r < r1; r += 8) {
      const float4 v = *reinterpret_cast<const float4*>(dy + (size_t)r * N + col);
      s.x += v.x, s.y += v.y, s.z += v.z, s.w += v.w;
    }
  part[wy][lane] = s;
  __syncthreads();
  if (wy == 0 && col < N) {
    float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int k = 0; k < 8; ++k) t.x += part[k][lane].x, t.y += part[k][lane].y, t.z += part[k][lane].z, t.w += part[k][lane].w;
    atomicAdd(&db[col], t.x);
    atomicAdd(&db[col + 1], t.y);
    atomicAdd(&db[col + 2], t.z);
    atomicAdd(&db[col + 3], t.w);
  }
}

// dx = dy where y > 0 else 0 (in place allowed).
__global__ void relu_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ y,
                                float* __restrict__ dx, size_t n) {
  pdl_wait();
  pdl_trigger();
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x)
    dx[i] = y[i] > 0.f ? dy[i] : 0.f;
}

// =======================================================================================
// Direct convolution fallbacks (any stride / zero padding, NHWC x HWIO).  Correct for every
// shape; used where the tensor-core implicit GEMM does not apply (e.g. C_in * 4 B < 16 B).
// =======================================================================================
struct ConvShape {
  int N, H, W, Ci, Co, KH, KW, OH, OW, sh, sw, ph, pw;
};

__global__ void conv2d_fwd_direct_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                         const float* __restrict__ bias, float* __restrict__ y,
                                         ConvShape s, int relu) {
  pdl_wait();
  pdl_trigger();
  const size_t total = (size_t)s.N * s.OH * s.OW * s.Co;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x) {
    const int co = i % s.Co;
    size_t t = i / s.Co;
    const int ow = t % s.OW;
    t /= s.OW;
    const int oh = t % s.OH;
    const int n = t / s.OH;
    float acc = bias != nullptr ? bias[co] : 0.f;
    for (int kh = 0; kh < s.KH; ++kh) {
      const int ih = oh * s.sh - s.ph + kh;
      if (ih < 0 || ih >= s.H) continue;
      for (int kw = 0; kw < s.KW; ++kw) {
        const int iw = ow * s.sw - s.pw + kw;
        if (iw < 0 || iw >= s.W) continue;
        const float* xp = x + (((size_t)n * s.H + ih) * s.W + iw) * s.Ci;
        const float* wp = w + ((size_t)(kh * s.KW + kw) * s.Ci) * s.Co + co;
        for (int ci = 0; ci < s.Ci; ++ci) acc = fmaf(xp[ci], wp[(size_t)ci * s.Co], acc);
      }
    }
    y[i] = relu ? fmaxf(acc, 0.f) : acc;
  }
}

// dx[n,ih,iw,ci] = sum_{kh,kw,co} dy[n,oh,ow,co] * w[kh,kw,ci,co]
__global__ void conv2d_dgrad_direct_kernel(const float* __restrict__ dy, const float* __restrict__ w,
                                           float* __restrict__ dx, ConvShape s) {
  pdl_wait();
  pdl_trigger();
  const size_t total = (size_t)s.N * s.H * s.W * s.Ci;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x) {
    const int ci = i % s.Ci;
    size_t t = i / s.Ci;
    const int iw = t % s.W;
    t /= s.W;
    const int ih = t % s.H;
    const int n = t / s.H;
    float acc = 0.f;
    for (int kh = 0; kh < s.KH; ++kh) {
      const int th = ih + s.ph - kh;
      if (th < 0 || th % s.sh) continue;
      const int oh = th / s.sh;
      if (oh >= s.OH) continue;
      for (int kw = 0; kw < s.KW; ++kw) {
        const int tw = iw + s.pw - kw;
        if (tw < 0 || tw % s.sw) continue;
        const int ow = tw / s.sw;
        if (ow >= s.OW) continue;
        const float* dp = dy + (((size_t)n * s.OH + oh) * s.OW + ow) * s.Co;
        const float* wp = w + ((size_t)(kh * s.KW + kw) * s.Ci + ci) * s.Co;
        for (int co = 0; co < s.Co; ++co) acc = fmaf(dp[co], wp[co], acc);
      }
    }
    dx[i] = acc;
  }
}

// dW[kh,kw,ci,co] += sum_{n,oh,ow} x[n,ih,iw,ci] * dy[n,oh,ow,co]; one CTA per (kh,kw,ci) x co-tile,
// threads stride over output pixels and reduce through shared memory.
__global__ void __launch_bounds__(256) conv2d_wgrad_direct_kernel(const float* __restrict__ x,
                                                                   const float* __restrict__ dy,
                                                                   float* __restrict__ dw,
                                                                   ConvShape s) {
  pdl_wait();
  pdl_trigger();
  __shared__ float part[8][33];
  const int tap = blockIdx.x;  // (kh*KW + kw)*Ci + ci
  const int ci = tap % s.Ci;
  const int kw = (tap / s.Ci) % s.KW;
  const int kh = tap / (s.Ci * s.KW);
  const int co = blockIdx.y * 32 + (threadIdx.x & 31);
  const int wy = threadIdx.x >> 5;
  const size_t npix = (size_t)s.N * s.OH * s.OW;
  float acc = 0.f;
  if (co < s.Co)
    for (size_t p = wy; p < npix; p += 8) {
      const int ow = p % s.OW;
      const int oh = (p / s.OW) % s.OH;
      const int n = p / ((size_t)s.OW * s.OH);
      const int ih = oh * s.sh - s.ph + kh, iw = ow * s.sw - s.pw + kw;
      if (ih < 0 || ih >= s.H || iw < 0 || iw >= s.W) continue;
      acc = fmaf(x[(((size_t)n * s.H + ih) * s.W + iw) * s.Ci + ci], dy[p * s.Co + co], acc);
    }
  part[wy][threadIdx.x & 31] = acc;
  __syncthreads();
  if (wy == 0 && co < s.Co) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += part[k][threadIdx.x];
    atomicAdd(&dw[(size_t)tap * s.Co + co], t);
  }
}

// =======================================================================================
// CUDA-core GEMM fallback with arbitrary strides (tiny or unaligned problems, e.g. N = 10).
//   D[m][n] (+)= relu?( sum_k A(m,k) B(k,n) + bias[n] )
// =======================================================================================
__global__ void __launch_bounds__(256) sgemm_simt_kernel(const float* __restrict__ A, long long sam,
                                                          long long sak, const float* __restrict__ Bm,
                                                          long long sbk, long long sbn,
                                                          float* __restrict__ D, long long ldd,
                                                          const float* __restrict__ bias, int M, int N,
                                                          int K, int relu, int accumulate) {
  pdl_wait();
  pdl_trigger();
  __shared__ float as[32][33], bs[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  const int m0 = blockIdx.y * 32, n0 = blockIdx.x * 32;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int k0 = 0; k0 < K; k0 += 32) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int rr = ty + r * 8;
      const int m = m0 + rr, k = k0 + tx;
      as[rr][tx] = (m < M && k < K) ? A[m * sam + k * sak] : 0.f;
      const int kk = k0 + rr, n = n0 + tx;
      bs[rr][tx] = (kk < K && n < N) ? Bm[kk * sbk + n * sbn] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 32; ++k) {
      const float bv = bs[k][tx];
#pragma unroll
      for (int r = 0; r < 4; ++r) acc[r] = fmaf(as[ty + r * 8][k], bv, acc[r]);
    }
    __syncthreads();
  }
  const int n = n0 + tx;
  if (n < N)
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int m = m0 + ty + r * 8;
      if (m < M) {
        float v = acc[r] + (bias != nullptr ? bias[n] : 0.f);
        if (relu) v = fmaxf(v, 0.f);
        if (accumulate)
          atomicAdd(&D[m * ldd + n], v);
        else
          D[m * ldd + n] = v;
      }
    }
}

// =======================================================================================
// Single-replica optimizer step: w -= lr * g ; g = 0 ; running metrics += step metrics. [K13]
// (With more than one replica this work is fused into the all-reduce kernels instead.)
// =======================================================================================
__global__ void sgd_apply_kernel(float* __restrict__ w, float* __restrict__ g, size_t n_params,
                                 size_t n_total, float lr) {
  pdl_wait();
  pdl_trigger();
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n_total;
       i += (size_t)gridDim.x * blockDim.x) {
    const float gv = g[i];
    w[i] = i < n_params ? w[i] - lr * gv : w[i] + gv;  // tail slots accumulate metrics
    g[i] = 0.f;
  }
}

__global__ void cast_f32_bf16_kernel(const float* __restrict__ src, uint16_t* __restrict__ dst,
                                     size_t n) {
  pdl_wait();
  pdl_trigger();
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x) {
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(0.f), "f"(src[i]));
    dst[i] = (uint16_t)(r & 0xffff);
  }
}

static inline int grid_for(size_t n, int block = 256, int ctas_per_sm = 8) {
  const int cap = sm_count() * ctas_per_sm;
  size_t g = (n + block - 1) / block;
  return (int)(g < 1 ? 1 : (g > (size_t)cap ? cap : g));
}

}  // namespace b2d

using namespace b2d;

// ---------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------
extern "C" int b2d_conv_c1_pool_fwd(const float* x, const float* w, const float* bias, float* y,
                                    uint8_t* code, int N, int H, int W, int Co, int KS,
                                    void* stream) {
  B2D_REQUIRE(KS == 3 || KS == 5, "conv_c1_pool: kernel size must be 3 or 5 (got %d)", KS);
  const int OH = H - KS + 1, OW = W - KS + 1, PH = OH / 2, PW = OW / 2;
  B2D_REQUIRE(PH > 0 && PW > 0, "conv_c1_pool: input too small");
  // Two bands per image at small batch keeps >= 128 CTAs in flight for batch 64.
  const int bands = N >= 2 * sm_count() ? 1 : (N >= sm_count() ? 2 : 4);
  const int band = ceil_div(PH, bands);
  const dim3 grid(N, ceil_div(PH, band));
  const size_t smem = (size_t)(2 * band + KS) * W * sizeof(float);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // README model (32 filters): one channel group, the small-footprint instantiation; wider layers reuse each
  // input window for two groups of 32 channels
  if (KS == 3 && Co <= 32)
    B2D_CUDA_CHECK(launch_pdl(conv_c1_pool_fwd_kernel<3, 1>, dim3(grid), dim3(256), smem, st, x, w, bias, y, code, H, W, Co, PH, PW, band));
  else if (KS == 3)
    B2D_CUDA_CHECK(launch_pdl(conv_c1_pool_fwd_kernel<3, 2>, dim3(grid), dim3(256), smem, st, x, w, bias, y, code, H, W, Co, PH, PW, band));
  else if (Co <= 32)
    B2D_CUDA_CHECK(launch_pdl(conv_c1_pool_fwd_kernel<5, 1>, dim3(grid), dim3(256), smem, st, x, w, bias, y, code, H, W, Co, PH, PW, band));
  else
    B2D_CUDA_CHECK(launch_pdl(conv_c1_pool_fwd_kernel<5, 2>, dim3(grid), dim3(256), smem, st, x, w, bias, y, code, H, W, Co, PH, PW, band));
  B2D_CUDA_CHECK(cudaGetLastError());
  return B2D_OK;
}

extern "C" int b2d_conv_c1_pool_bwd(const float* x, const float* dy, const uint8_t* code, float* dw,
                                    float* db, int N, int H, int W, int Co, int KS, void* stream) {
  B2D_REQUIRE(KS == 3 || KS == 5, "conv_c1_pool: kernel size must be 3 or 5 (got %d)", KS);
  const int OH = H - KS + 1, OW = W - KS + 1, PH = OH / 2, PW = OW / 2;
  // Backward pays a cross-warp reduction + ~10*Co global atomics per CTA: fewer, fatter CTAs.
  const int bands = N >= 120 ? 1 : 2;
  const int band = ceil_div(PH, bands);
  const dim3 grid(N, ceil_div(PH, band));
  const size_t smem = ((size_t)(2 * band + KS) * W + 8 * (KS * KS + 1) * 32) * sizeof(float);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (KS == 3)
    B2D_CUDA_CHECK(launch_pdl(conv_c1_pool_bwd_kernel<3>, dim3(grid), dim3(256), smem, st, x, dy, code, dw, db, H, W, Co, PH, PW, band));
  else
    B2D_CUDA_CHECK(launch_pdl(conv_c1_pool_bwd_kernel<5>, dim3(grid), dim3(256), smem, st, x, dy, code, dw, db, H, W, Co, PH, PW, band));
  B2D_CUDA_CHECK(cudaGetLastError());
  return B2D_OK;
}

extern "C" int b2d_dense_xent_head(const float* h, const float* w, const float* bias,
                                   const int* labels, float* logits_out, float* dh, float* dw,
                                   float* db, float* dh_colsum, float* metrics, int B, int K, int C,
                                   float grad_scale, int relu_mask, void* stream) {
  B2D_REQUIRE(C >= 1 && C <= kHeadMaxC, "dense_xent_head: C must be in [1, %d]", kHeadMaxC);
  // Large batches: the throughput-oriented variant (16 classes at most, K * 16 / 256 outputs per thread <= 64)
  if (B >= 256 && C <= 16 && K <= 1024 && K % 32 == 0) {
    const size_t smem_w = ((size_t)K * 20 + 16 * K + 8 * 16 + 16) * sizeof(float);
    static bool attr_w = false;
    if (!attr_w) {
      B2D_CUDA_CHECK(cudaFuncSetAttribute(dense_xent_head_wide_kernel<16>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
      attr_w = true;
    }
    // 8 rows per CTA until that is more than two waves of CTAs; then 16 (half the global atomics)
    const int passes = B > 8 * 2 * sm_count() ? 2 : 1;
    B2D_CUDA_CHECK(launch_pdl(dense_xent_head_wide_kernel<16>, dim3(ceil_div(B, 8 * passes)), dim3(256), smem_w,
                              static_cast<cudaStream_t>(stream), h, w, bias, labels, logits_out, dh, dw, db,
                              dh_colsum, metrics, B, K, C, grad_scale, relu_mask, passes));
    B2D_CUDA_CHECK(cudaGetLastError());
    return B2D_OK;
  }
  const size_t smem =
      ((size_t)2 * K * C + (2 * kHeadRows + 1) * K + kHeadRows * kHeadMaxC + 2 * kHeadRows) * sizeof(float);
  B2D_REQUIRE(smem <= 200 * 1024, "dense_xent_head: K*C too large for the fused head (K=%d C=%d)", K, C);
  static bool attr_set = false;
  if (!attr_set) {
    B2D_CUDA_CHECK(cudaFuncSetAttribute(dense_xent_head_kernel,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr_set = true;
  }
  // One pass of 8 rows per CTA at small batch; several passes per CTA once the grid is full.
  // Every CTA ends with K*C global atomics onto the same addresses: keep the CTA count modest
  // (<= 64) once the batch is large, trading a little parallelism for 2-4x fewer atomics.
  int rows = kHeadRows;
  if (B > kHeadRows * 64) rows = ceil_div(ceil_div(B, 64), kHeadRows) * kHeadRows;
  B2D_CUDA_CHECK(launch_pdl(dense_xent_head_kernel, dim3(ceil_div(B, rows)), dim3(256), smem, static_cast<cudaStream_t>(stream), h, w, bias, labels, logits_out, dh, dw, db, dh_colsum, metrics, B, K, C, grad_scale, relu_mask, rows));
  B2D_CUDA_CHECK(cudaGetLastError());
  return B2D_OK;
}

extern "C" int b2d_softmax_xent(const float* logits, const int* labels, float* dlogits,
                                float* metrics, int B, int C, float grad_scale, void* stream) {
  B2D_CUDA_CHECK(launch_pdl(softmax_xent_kernel, dim3(ceil_div(B * 32, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), logits, labels, dlogits, metrics, B, C, grad_scale));
  B2D_CUDA_CHECK(cudaGetLastError());
  return B2D_OK;
}

extern "C" int b2d_maxpool2_fwd(const float* x, float* y, uint8_t* code, int N, int H, int W, int C,
                                void* stream) {
  const int PH = H / 2, PW = W / 2;
  B2D_CUDA_CHECK(launch_pdl(maxpool2_fwd_kernel, dim3(grid_for((size_t)N * PH * PW * C)), dim3(256), 0, static_cast<cudaStream_t>(stream), x, y, code, N, H, W, C, PH, PW));
  B2D_CUDA_CHECK(cudaGetLastError());
  return B2D_OK;
}

extern "C" int b2d_maxpool2_bwd(const float* dy, const uint8_t* code, const float* y_pooled, float* dx,
                                int N, int H, int W, int C, int act_relu, void* stream) {
  const int PH = H / 2, PW = W / 2;
  if (C % 4 == 0) {
    const size_t n4 = (size_t)N * H * W * (C / 4);
    B2D_CUDA_CHECK(launch_pdl(maxpool2_bwd_v4_kernel, dim3(grid_for(n4, 256, 32)), dim3(256), 0,
                              static_cast<cudaStream_t>(stream), reinterpret_cast<const float4*>(dy),
                              reinterpret_cast<const uchar4*>(code), reinterpret_cast<const float4*>(y_pooled),
                              reinterpret_cast<float4*>(dx), N, H, W, C / 4, PH, PW, act_relu));
    return B2D_OK;
  }
  B2D_CUDA_CHECK(launch_pdl(maxpool2_bwd_kernel, dim3(grid_for((size_t)N * H * W * C)), dim3(256), 0, static_cast<cudaStream_t>(stream), dy, code, y_pooled, dx, N, H, W, C, PH, PW, act_relu));
  B2D_CUDA_CHECK(cudaGetLastError());
  return B2D_OK;
}

extern "C" int b2d_colsum(const float* dy, float* db, int M, int N, void* stream) {
  if (N % 4 == 0 && M >= 1024) {
    const int rows_v4 = max(64, ceil_div(M, 512));
    B2D_CUDA_CHECK(launch_pdl(colsum_v4_kernel, dim3(ceil_div(N, 128), ceil_div(M, rows_v4)), dim3(256), 0,
                              static_cast<cudaStream_t>(stream), dy, db, M, N, rows_v4));
    return B2D_OK;
  }
  const int rows = max(8, ceil_div(M, 64));
  const dim3 grid(ceil_div(N, 32), ceil_div(M, rows));
  B2D_CUDA_CHECK(launch_pdl(colsum_kernel, dim3(grid), dim3(256), 0, static_cast<cudaStream_t>(stream), dy, db, M, N, rows));
  B2D_CUDA_CHECK(cudaGetLastError());
  return B2D_OK;
}

extern "C" int b2d_relu_bwd(const float* dy, const float* y, float* dx, size_t n, void* stream) {
  B2D_CUDA_CHECK(launch_pdl(relu_bwd_kernel, dim3(grid_for(n)), dim3(256), 0, static_cast<cudaStream_t>(stream), dy, y, dx, n));
  B2D_CUDA_CHECK(cudaGetLastError());
  return B2D_OK;
}

// ph / pw are the padding at the top / left; the bottom / right padding is whatever the given
// output size implies (TensorFlow's 'same' puts the odd pixel at the end), since reads outside
// the image are skipped.  OH / OW <= 0 mean "derive from symmetric padding".
static ConvShape make_shape(int N, int H, int W, int Ci, int Co, int KH, int KW, int sh, int sw,
                            int ph, int pw, int OH, int OW) {
  if (OH <= 0) OH = (H + 2 * ph - KH) / sh + 1;
  if (OW <= 0) OW = (W + 2 * pw - KW) / sw + 1;
  ConvShape s{N, H, W, Ci, Co, KH, KW, OH, OW, sh, sw, ph, pw};
  return s;
}

extern "C" int b2d_conv2d_fwd_direct(const float* x, const float* w, const float* bias, float* y,
                                     int N, int H, int W, int Ci, int Co, int KH, int KW, int sh,
                                     int sw, int ph, int pw, int OH, int OW, int relu, void* stream) {
  const ConvShape s = make_shape(N, H, W, Ci, Co, KH, KW, sh, sw, ph, pw, OH, OW);
  B2D_CUDA_CHECK(launch_pdl(conv2d_fwd_direct_kernel, dim3(grid_for((size_t)N * s.OH * s.OW * Co)), dim3(256), 0, static_cast<cudaStream_t>(stream), x, w, bias, y, s, relu));
  B2D_CUDA_CHECK(cudaGetLastError());
  return B2D_OK;
}

extern "C" int b2d_conv2d_dgrad_direct(const float* dy, const float* w, float* dx, int N, int H, int W,
                                       int Ci, int Co, int KH, int KW, int sh, int sw, int ph, int pw,
                                       int OH, int OW, void* stream) {
  const ConvShape s = make_shape(N, H, W, Ci, Co, KH, KW, sh, sw, ph, pw, OH, OW);
  B2D_CUDA_CHECK(launch_pdl(conv2d_dgrad_direct_kernel, dim3(grid_for((size_t)N * H * W * Ci)), dim3(256), 0, static_cast<cudaStream_t>(stream), dy, w, dx, s));
  B2D_CUDA_CHECK(cudaGetLastError());
  return B2D_OK;
}

extern "C" int b2d_conv2d_wgrad_direct(const float* x, const float* dy, float* dw, int N, int H, int W,
                                       int Ci, int Co, int KH, int KW, int sh, int sw, int ph, int pw,
                                       int OH, int OW, void* stream) {
  const ConvShape s = make_shape(N, H, W, Ci, Co, KH, KW, sh, sw, ph, pw, OH, OW);
  const dim3 grid(KH * KW * Ci, ceil_div(Co, 32));
  B2D_CUDA_CHECK(launch_pdl(conv2d_wgrad_direct_kernel, dim3(grid), dim3(256), 0, static_cast<cudaStream_t>(stream), x, dy, dw, s));
  B2D_CUDA_CHECK(cudaGetLastError());
  return B2D_OK;
}

extern "C" int b2d_sgemm_simt(const float* A, int64_t sam, int64_t sak, const float* B, int64_t sbk,
                              int64_t sbn, float* D, int64_t ldd, const float* bias, int M, int N,
                              int K, int relu, int accumulate, void* stream) {
  const dim3 grid(ceil_div(N, 32), ceil_div(M, 32));
  B2D_CUDA_CHECK(launch_pdl(sgemm_simt_kernel, dim3(grid), dim3(256), 0, static_cast<cudaStream_t>(stream), A, sam, sak, B, sbk, sbn, D, ldd, bias, M, N, K, relu, accumulate));
  B2D_CUDA_CHECK(cudaGetLastError());
  return B2D_OK;
}

extern "C" int b2d_sgd_apply(float* w, float* g, size_t n_params, size_t n_total, float lr,
                             void* stream) {
  B2D_CUDA_CHECK(launch_pdl(sgd_apply_kernel, dim3(grid_for(n_total)), dim3(256), 0, static_cast<cudaStream_t>(stream), w, g, n_params, n_total, lr));
  B2D_CUDA_CHECK(cudaGetLastError());
  return B2D_OK;
}

extern "C" int b2d_cast_f32_bf16(const float* src, void* dst, size_t n, void* stream) {
  B2D_CUDA_CHECK(launch_pdl(cast_f32_bf16_kernel, dim3(grid_for(n)), dim3(256), 0, static_cast<cudaStream_t>(stream), src, static_cast<uint16_t*>(dst), n));
  B2D_CUDA_CHECK(cudaGetLastError());
  return B2D_OK;
}
