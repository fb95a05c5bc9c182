// gemm_tcgen05.cu — the tensor-core GEMM core of distributed_b200 (sm_100a only).
//
// One CTA owns one 128 x BLOCK_N output tile over one K slice:
//   warp 0   TMA producer: cp.async.bulk.tensor tiles (128B-swizzled) into a STAGES-deep ring
//   warp 1   MMA issuer:   one thread issues tcgen05.mma (kind::tf32 / kind::f16), fp32
//                          accumulator in tensor memory; tcgen05.commit frees ring slots
//   warp 2   TMEM allocator / deallocator
//   warps 4-7 epilogue:    tcgen05.ld -> bias / ReLU / ReLU-mask -> store, atomic accumulate,
//                          or split-K meet-in-workspace with a last-CTA finalize
// Both operands can be K-major or MN-major, which covers Dense forward (x.W), dgrad (dy.W^T)
// and wgrad (x^T.dy) on Keras' (in, out) kernel layout without any transposition pass.
//
// Reference behaviour: Dense layers of the README model (README.md:296-297); the reference
// itself ships no kernels.
#include <cuda_bf16.h>

#include <cstdlib>
#include <new>

#include "host_common.h"
#include "ptx.cuh"

namespace b2d {

constexpr int kBlockM = 128;
constexpr int kABytes = kBlockM * 128;  // one stage of A: 128 rows (or 128 MN elements) x 128 B
constexpr int kSmemBudget = 196 * 1024;
constexpr int kMaxStages = 8;

enum GemmFlags : int {
  kFlagRelu = 1,
  kFlagAccumulate = 2,  // D += tile via fp32 vector atomics
  kFlagFinalize = 4,    // split-K: meet in workspace, last CTA applies the epilogue
  kFlagOutBf16 = 8,
  kFlagMaskBf16 = 16,   // the ReLU mask tensor is bf16 (same row pitch convention, in elements)
  kFlagScatter = 32,    // accumulate into the gradient arena of the replica that OWNS each element (reduce-scatter)
};

// GEMM -> reduce-scatter in one kernel: D is a range of the replicated gradient arena, cut into `world`
// equal shards; every 16-byte vector of the result is added (red.add over NVLink) into the arena of the
// replica that owns its shard, so when all replicas' GEMMs have finished each owner holds the complete sum
// of its shard and nobody has to fetch anything (csrc/cuda/allreduce.cu, kScattered).
struct ScatterArgs {
  float* base[8];        // gradient arena of every replica (peer-mapped); base[rank] is local
  long long d_off;       // float offset of D[0][0] inside the arena
  long long seg_begin;   // float offset of the sharded range
  long long per;         // floats per shard (multiple of 4)
  int world;
};

// Implicit-GEMM convolution view of the A operand (stride-1 NHWC convolutions): the M tile is a
// box (n_t images x ho_t rows x wo_t columns) of output pixels, and k-block kb = (tap, channel
// chunk) loads the input box shifted by that filter tap with ONE 4-D TMA copy — out-of-range
// pixels (zero padding) are filled by the TMA unit.  No im2col buffer exists anywhere.
struct ConvGeom {
  int enabled;
  int kw, c_chunks;        // filter width; 128-byte channel chunks per tap
  int wo_t, ho_t, n_t;     // output-pixel box of one M tile (wo_t * ho_t * n_t <= 128 rows)
  int rows_used;           // wo_t * ho_t * n_t
  int OH, OW, N_img;       // output geometry
  int tiles_h;             // M tiles per image (when n_t == 1)
  int w_base, h_base;      // input coordinate of tap (0,0) for output pixel (0,0)
  int tap_sign;            // +1 forward (in = out + tap - pad), -1 dgrad (in = out + pad - tap)
  int b_rows_per_tap;      // K-major B (dgrad): rows of the weight matrix per tap (= C_in)
  int stride;              // forward only: input step per output pixel (TMA element stride)
  // strided data gradient, one launch per output-pixel parity class: the class's taps are
  // r = tap_r0 + tap_step * tr (same for columns) of a kw_full-wide filter, and its output pixels
  // are (out_step * oh + out_off_h, out_step * ow + out_off_w) of an OH_full x OW_full image.
  int tap_r0, tap_s0, tap_step, kw_full;
  int out_step, out_off_h, out_off_w, OH_full, OW_full;
};

struct GemmArgs {
  CUtensorMap tmA;
  CUtensorMap tmB;
  ConvGeom cv;
  long long* dbg;  // optional: per-CTA timeline (globaltimer ns) for latency analysis, 16 slots per CTA
  int stages;
  void* D;
  const float* bias;
  const void* mask;
  const void* addend;  // optional: D = epilogue(acc + addend), same dtype / row pitch as D (gradient sums)
  float* ws;
  uint32_t* counters;
  long long ldd;
  long long ldmask;
  long long ldws;
  int M, N, K;
  int kb_total, kb_per_split, splits;
  int m_tiles, n_tiles;   // tile grid (persistent CTAs walk m_tiles * n_tiles * splits tiles)
  int flags;
  ScatterArgs sc;
};

__device__ __forceinline__ long long gtime_ns() {
  long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ void red_add_f4(float* p, float a, float b, float c, float d) {
  asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b),
               "f"(c), "f"(d)
               : "memory");
}

__device__ __forceinline__ void red_add_f4_sys(float* p, float a, float b, float c, float d) {
  asm volatile("red.relaxed.sys.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b),
               "f"(c), "f"(d)
               : "memory");
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

// Applies bias / ReLU / mask to four consecutive columns of one row and stores them.
template <bool OUT_BF16, bool SCATTER = false>
__device__ __forceinline__ void epilogue_store4(const GemmArgs& a, int row, int col, float v0,
                                                float v1, float v2, float v3) {
  if (a.bias != nullptr) {
    const float4 b = __ldg(reinterpret_cast<const float4*>(a.bias + col));
    v0 += b.x, v1 += b.y, v2 += b.z, v3 += b.w;
  }
  if (a.flags & kFlagRelu) {
    v0 = fmaxf(v0, 0.f), v1 = fmaxf(v1, 0.f), v2 = fmaxf(v2, 0.f), v3 = fmaxf(v3, 0.f);
  }
  if (a.addend != nullptr) {
    if (OUT_BF16 && (a.flags & kFlagOutBf16)) {
      const uint2 t = *reinterpret_cast<const uint2*>(static_cast<const uint16_t*>(a.addend) + (long long)row * a.ldd + col);
      v0 += __uint_as_float(t.x << 16), v1 += __uint_as_float(t.x & 0xffff0000u);
      v2 += __uint_as_float(t.y << 16), v3 += __uint_as_float(t.y & 0xffff0000u);
    } else {
      const float4 t = __ldg(reinterpret_cast<const float4*>(static_cast<const float*>(a.addend) + (long long)row * a.ldd + col));
      v0 += t.x, v1 += t.y, v2 += t.z, v3 += t.w;
    }
  }
  if (a.mask != nullptr) {
    float4 m;
    if (a.flags & kFlagMaskBf16) {
      const uint2 t = *reinterpret_cast<const uint2*>(static_cast<const uint16_t*>(a.mask) +
                                                      (long long)row * a.ldmask + col);
      m = make_float4(__uint_as_float(t.x << 16), __uint_as_float(t.x & 0xffff0000u),
                      __uint_as_float(t.y << 16), __uint_as_float(t.y & 0xffff0000u));
    } else {
      m = __ldg(reinterpret_cast<const float4*>(static_cast<const float*>(a.mask) + (long long)row * a.ldmask + col));
    }
    v0 = m.x > 0.f ? v0 : 0.f, v1 = m.y > 0.f ? v1 : 0.f;
    v2 = m.z > 0.f ? v2 : 0.f, v3 = m.w > 0.f ? v3 : 0.f;
  }
  if (SCATTER) {      // compiled into the reduce-scatter instantiations only: run-once kernels pay for every instruction
    const long long idx = a.sc.d_off + (long long)row * a.ldd + col;
    const int owner = min(a.sc.world - 1, static_cast<int>((idx - a.sc.seg_begin) / a.sc.per));
    red_add_f4_sys(a.sc.base[owner] + idx, v0, v1, v2, v3);
  } else if (a.flags & kFlagAccumulate) {
    red_add_f4(reinterpret_cast<float*>(a.D) + (long long)row * a.ldd + col, v0, v1, v2, v3);
  } else if (OUT_BF16 && (a.flags & kFlagOutBf16)) {
    uint2 p;
    p.x = pack_bf16x2(v0, v1);
    p.y = pack_bf16x2(v2, v3);
    *reinterpret_cast<uint2*>(reinterpret_cast<uint16_t*>(a.D) + (long long)row * a.ldd + col) = p;
  } else {
    *reinterpret_cast<float4*>(reinterpret_cast<float*>(a.D) + (long long)row * a.ldd + col) =
        make_float4(v0, v1, v2, v3);
  }
}

// bf16 output, eight consecutive columns of one row as ONE 16-byte store (plain store only: no
// accumulate / split-K).  Requires col + 8 <= N and 16-byte aligned rows (ldd % 8 == 0).
__device__ __forceinline__ void epilogue_store8_bf16(const GemmArgs& a, int row, int col, float (&v)[8]) {
  if (a.bias != nullptr) {
    const float4 b0 = *reinterpret_cast<const float4*>(a.bias + col);
    const float4 b1 = *reinterpret_cast<const float4*>(a.bias + col + 4);
    v[0] += b0.x, v[1] += b0.y, v[2] += b0.z, v[3] += b0.w;
    v[4] += b1.x, v[5] += b1.y, v[6] += b1.z, v[7] += b1.w;
  }
  if (a.flags & kFlagRelu) {
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = fmaxf(v[k], 0.f);
  }
  if (a.addend != nullptr) {
    const uint4 t = *reinterpret_cast<const uint4*>(static_cast<const uint16_t*>(a.addend) + (long long)row * a.ldd + col);
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) v[2 * k] += __uint_as_float(w[k] << 16), v[2 * k + 1] += __uint_as_float(w[k] & 0xffff0000u);
  }
  if (a.mask != nullptr) {
    float m[8];
    if (a.flags & kFlagMaskBf16) {
      const uint4 t = *reinterpret_cast<const uint4*>(static_cast<const uint16_t*>(a.mask) +
                                                      (long long)row * a.ldmask + col);
      const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) m[2 * k] = __uint_as_float(w[k] << 16), m[2 * k + 1] = __uint_as_float(w[k] & 0xffff0000u);
    } else {
      const float* mp = static_cast<const float*>(a.mask) + (long long)row * a.ldmask + col;
      const float4 m0 = *reinterpret_cast<const float4*>(mp), m1 = *reinterpret_cast<const float4*>(mp + 4);
      m[0] = m0.x, m[1] = m0.y, m[2] = m0.z, m[3] = m0.w, m[4] = m1.x, m[5] = m1.y, m[6] = m1.z, m[7] = m1.w;
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = m[k] > 0.f ? v[k] : 0.f;
  }
  uint4 p;
  p.x = pack_bf16x2(v[0], v[1]), p.y = pack_bf16x2(v[2], v[3]);
  p.z = pack_bf16x2(v[4], v[5]), p.w = pack_bf16x2(v[6], v[7]);
  *reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(a.D) + (long long)row * a.ldd + col) = p;
}

template <MmaKind KIND, bool A_MN, bool B_MN, int BLOCK_N>
struct GemmCfg {
  static constexpr int ES = (KIND == MmaKind::TF32) ? 4 : 2;  // operand element size
  static constexpr int BLOCK_K = 128 / ES;                    // elements per k-block (one 128 B row)
  static constexpr int UMMA_K = 32 / ES;                      // K per tcgen05.mma
  static constexpr int ATOM_E = 128 / ES;                     // MN elements per 128 B swizzle atom
  static constexpr int B_BYTES = BLOCK_N * 128;
  static constexpr int STAGE_BYTES = kABytes + B_BYTES;
  static constexpr int STAGES =
      (kSmemBudget / STAGE_BYTES) < kMaxStages ? (kSmemBudget / STAGE_BYTES) : kMaxStages;
  static constexpr int SMEM_BYTES =
      STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
  static constexpr uint32_t TMEM_COLS = BLOCK_N < 32 ? 32 : BLOCK_N;
  // MN-major tiles: 16-bit operands use the plain 128 B swizzle (8-row atoms, 1024 B);
  // 32-bit operands must use the 32 B-chunk variant (4-row atoms, 512 B).
  static constexpr uint32_t MN_LAYOUT = (KIND == MmaKind::TF32) ? kUmmaSw128Base32 : kUmmaSw128;
  static constexpr uint32_t MN_SBO = (KIND == MmaKind::TF32) ? 512 : 1024;
  static constexpr uint32_t IDESC =
      umma_idesc(KIND == MmaKind::TF32 ? 2u : 1u, A_MN ? 1u : 0u, B_MN ? 1u : 0u, kBlockM, BLOCK_N);
};

template <MmaKind KIND, bool A_MN, bool B_MN, int BLOCK_N, bool OUT_BF16, bool SCATTER = false>
__global__ void __launch_bounds__(256, 2) gemm_tc_kernel(const __grid_constant__ GemmArgs args) {
  // Persistent CTA: tiles (m, n, k-split) are walked with stride gridDim.x.  The accumulator is
  // double-buffered in tensor memory, so the epilogue warps drain tile t while the MMA thread is already
  // filling tile t+1, and the TMA producer runs ahead across tile boundaries: the per-CTA prologue is paid
  // once per SM and the epilogue of every tile but the last is hidden.
  using Cfg = GemmCfg<KIND, A_MN, B_MN, BLOCK_N>;
  const int STAGES = args.stages;  // ring depth chosen on the host (<= Cfg::STAGES)
  constexpr int STAGE_BYTES = Cfg::STAGE_BYTES;
  constexpr int BLOCK_K = Cfg::BLOCK_K;
  constexpr uint32_t ACC_COLS = Cfg::TMEM_COLS;      // one accumulator

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;      // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;      // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);
  int* last_flag = reinterpret_cast<int*>(tmem_slot + 1);

  const int warp = threadIdx.x >> 5;  // warp-uniform
  const int lane = threadIdx.x & 31;
  long long* dbg = args.dbg == nullptr ? nullptr : args.dbg + 16 * blockIdx.x;
  if (dbg != nullptr && threadIdx.x == 0) dbg[0] = gtime_ns();
  const ConvGeom& cv = args.cv;
  const int total_tiles = args.m_tiles * args.n_tiles * args.splits;
  // two accumulators only when this CTA walks more than one tile (a one-tile CTA has nothing to overlap)
  const uint32_t tmem_cols = total_tiles > static_cast<int>(gridDim.x) ? 2 * ACC_COLS : ACC_COLS;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&args.tmA);
    tma_prefetch_desc(&args.tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tmem_full_bar[b], 1);
      mbar_init(&tmem_empty_bar[b], 4);   // one arrival per epilogue warp
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc_dyn(tmem_slot, tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // Everything above (barrier init, TMEM allocation, descriptor prefetch) overlapped with the
  // tail of the previous kernel; from here on its results are needed.
  if (dbg != nullptr && threadIdx.x == 0) dbg[1] = gtime_ns();   // prologue done
  pdl_wait();
  pdl_trigger();
  if (dbg != nullptr && threadIdx.x == 0) dbg[2] = gtime_ns();   // upstream kernel complete

  // tile id -> (m tile, n tile, k split): n fastest, so that neighbouring CTAs share the A tile in L2
  auto decode = [&](int tile, int& mt, int& nt, int& kb_begin, int& num_kb) {
    nt = tile % args.n_tiles;
    const int r = tile / args.n_tiles;
    mt = r % args.m_tiles;
    const int z = r / args.m_tiles;
    kb_begin = z * args.kb_per_split;
    num_kb = min(args.kb_total, kb_begin + args.kb_per_split) - kb_begin;   // >= 1 by construction
  };
  // convolution view: first image / first output row of an M tile
  auto conv_origin = [&](int mt, int& cv_n0, int& cv_oh0) {
    cv_n0 = 0, cv_oh0 = 0;
    if (cv.enabled) {
      if (cv.n_t == 1) {
        cv_n0 = mt / cv.tiles_h;
        cv_oh0 = (mt % cv.tiles_h) * cv.ho_t;
      } else {
        cv_n0 = mt * cv.n_t;
      }
    }
  };

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      uint32_t it = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        int mt, nt, kb_begin, num_kb, cv_n0, cv_oh0;
        decode(tile, mt, nt, kb_begin, num_kb);
        conv_origin(mt, cv_n0, cv_oh0);
        const int m0 = mt * kBlockM, n0 = nt * BLOCK_N;
        for (int i = 0; i < num_kb; ++i, ++it) {
          const int stage = it % STAGES;
          const uint32_t phase = (it / STAGES) & 1;
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * STAGE_BYTES;
          uint8_t* sb = sa + kABytes;
          const int kb = kb_begin + i;
          const int k0 = kb * BLOCK_K;
          int tap = 0, cc = 0;
          if (cv.enabled) tap = kb / cv.c_chunks, cc = kb - tap * cv.c_chunks;
          mbar_expect_tx(&full_bar[stage], cv.enabled ? cv.rows_used * 128 + Cfg::B_BYTES : STAGE_BYTES);
          if constexpr (!A_MN) {
            if (cv.enabled) {
              const int r = tap / cv.kw, sx = tap - r * cv.kw;
              tma_load_4d(sa, &args.tmA, &full_bar[stage], cc * BLOCK_K, cv.w_base + cv.tap_sign * sx,
                          cv_oh0 * cv.stride + cv.h_base + cv.tap_sign * r, cv_n0);
            } else {
              tma_load_2d(sa, &args.tmA, &full_bar[stage], k0, m0);
            }
          } else {
#pragma unroll
            for (int a = 0; a < kBlockM / Cfg::ATOM_E; ++a)
              tma_load_2d(sa + a * (BLOCK_K * 128), &args.tmA, &full_bar[stage], m0 + a * Cfg::ATOM_E, k0);
          }
          if constexpr (!B_MN) {
            if (cv.enabled) {  // dgrad: weights [taps * C_in rows][C_out], K = C_out chunk of this tap
              const int tr = tap / cv.kw, ts = tap - tr * cv.kw;
              const int tap_full = (cv.tap_r0 + cv.tap_step * tr) * cv.kw_full + cv.tap_s0 + cv.tap_step * ts;
              tma_load_2d(sb, &args.tmB, &full_bar[stage], cc * BLOCK_K, tap_full * cv.b_rows_per_tap + n0);
            } else {
              tma_load_2d(sb, &args.tmB, &full_bar[stage], k0, n0);
            }
          } else {
#pragma unroll
            for (int a = 0; a < (BLOCK_N + Cfg::ATOM_E - 1) / Cfg::ATOM_E; ++a)
              tma_load_2d(sb + a * (BLOCK_K * 128), &args.tmB, &full_bar[stage], n0 + a * Cfg::ATOM_E, k0);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (single thread) =====
    if (lane == 0) {
      uint32_t it = 0, t = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++t) {
        int mt, nt, kb_begin, num_kb;
        decode(tile, mt, nt, kb_begin, num_kb);
        const uint32_t buf = t & 1;
        mbar_wait(&tmem_empty_bar[buf], ((t >> 1) & 1) ^ 1);   // epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t acc = tmem_base + buf * ACC_COLS;
        for (int i = 0; i < num_kb; ++i, ++it) {
          const int stage = it % STAGES;
          const uint32_t phase = (it / STAGES) & 1;
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (dbg != nullptr && it == 0) dbg[3] = gtime_ns();           // first tiles landed
          const uint32_t a_base = smem_u32(smem + stage * STAGE_BYTES);
          const uint32_t b_base = a_base + kABytes;
#pragma unroll
          for (int j = 0; j < BLOCK_K / Cfg::UMMA_K; ++j) {
            // K-major: step 32 B inside the 128 B swizzled row; 8-row groups are 1024 B apart.
            // MN-major: step UMMA_K rows (whole swizzle atoms); LBO = distance between the
            // 128 B-wide MN atoms (one TMA box each), SBO = one atom of K rows (8, or 4 for tf32).
            const uint64_t da = A_MN ? umma_smem_desc_sw128(a_base + j * Cfg::UMMA_K * 128,
                                                            BLOCK_K * 128, Cfg::MN_SBO, Cfg::MN_LAYOUT)
                                     : umma_smem_desc_sw128(a_base + j * 32, 16, 1024);
            const uint64_t db = B_MN ? umma_smem_desc_sw128(b_base + j * Cfg::UMMA_K * 128,
                                                            BLOCK_K * 128, Cfg::MN_SBO, Cfg::MN_LAYOUT)
                                     : umma_smem_desc_sw128(b_base + j * 32, 16, 1024);
            tc_mma<KIND>(acc, da, db, Cfg::IDESC, (i | j) != 0 ? 1u : 0u);
          }
          tc_commit(&empty_bar[stage]);  // frees this ring slot once the MMAs above retire
        }
        tc_commit(&tmem_full_bar[buf]);  // accumulator complete
        if (dbg != nullptr && t == 0) dbg[4] = gtime_ns();             // all MMAs of the first tile issued
      }
    }
  } else if (warp >= 4) {
    // ===== epilogue: 4 warps, warp q owns TMEM lanes [32q, 32q+32) =====
    const int q = warp - 4;
    const bool finalize = (args.flags & kFlagFinalize) != 0;
    // bf16 output written directly (no split-K meeting, no accumulation): 16-byte stores
    const bool wide_bf16 = OUT_BF16 &&
                           (args.flags & (kFlagOutBf16 | kFlagFinalize | kFlagAccumulate)) == kFlagOutBf16 &&
                           (args.ldd & 7) == 0 && (args.mask == nullptr || (args.ldmask & 7) == 0);
    uint32_t t = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++t) {
      int mt, nt, kb_begin, num_kb, cv_n0, cv_oh0;
      decode(tile, mt, nt, kb_begin, num_kb);
      conv_origin(mt, cv_n0, cv_oh0);
      const int m0 = mt * kBlockM, n0 = nt * BLOCK_N;
      const uint32_t buf = t & 1;
      int row = m0 + q * 32 + lane;
      bool row_ok = row < args.M;
      if (cv.enabled) {  // tile-local pixel index -> (image, out row, out column) -> output matrix row
        const int m = q * 32 + lane;
        const int w_i = m % cv.wo_t, tt = m / cv.wo_t;
        const int h_i = tt % cv.ho_t, n_i = tt / cv.ho_t;
        const int oh = cv_oh0 + h_i, n = cv_n0 + n_i;
        row_ok = m < cv.rows_used && oh < cv.OH && n < cv.N_img;
        row = (n * cv.OH_full + oh * cv.out_step + cv.out_off_h) * cv.OW_full + w_i * cv.out_step + cv.out_off_w;
      }
      // Fast path of the bf16 networks: the epilogue operands that do NOT depend on the accumulator (the
      // other branch's gradient, the ReLU mask) are fetched one 32-column chunk ahead — the first chunk
      // while the main loop is still running — as read-only loads the compiler cannot serialise behind the
      // stores to D.
      const bool fast = wide_bf16 && row_ok && (args.mask == nullptr || (args.flags & kFlagMaskBf16));
      const uint16_t* add_row =
          args.addend == nullptr ? nullptr : static_cast<const uint16_t*>(args.addend) + (long long)row * args.ldd;
      const uint16_t* mask_row =
          args.mask == nullptr ? nullptr : static_cast<const uint16_t*>(args.mask) + (long long)row * args.ldmask;
      uint4 pa[4], pm[4];
      auto prefetch = [&](int c) {
        const int col0 = n0 + c * 32;
        if (fast && col0 + 32 <= args.N) {
          if (add_row != nullptr) {
#pragma unroll
            for (int g = 0; g < 4; ++g) pa[g] = __ldg(reinterpret_cast<const uint4*>(add_row + col0) + g);
          }
          if (mask_row != nullptr) {
#pragma unroll
            for (int g = 0; g < 4; ++g) pm[g] = __ldg(reinterpret_cast<const uint4*>(mask_row + col0) + g);
          }
        }
      };
      if constexpr (OUT_BF16) prefetch(0);
      mbar_wait(&tmem_full_bar[buf], (t >> 1) & 1);
      tc_fence_after();
      if (dbg != nullptr && threadIdx.x == 128 && t == 0) dbg[5] = gtime_ns();   // accumulator visible
      const uint32_t acc = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * ACC_COLS;
      // Each thread owns one accumulator row (TMEM lane) and stores 32 consecutive columns per chunk.
      // (A shared-memory transposed, fully coalesced variant measured the same latency on the small
      // GEMMs and 9% lower throughput at 8192^3, so the direct form stays.)
#pragma unroll 1
      for (int c = 0; c < BLOCK_N / 32; ++c) {
        uint32_t v[32];
        tmem_ld_32x32(acc + c * 32, v);
        tmem_ld_wait();
        if (c == BLOCK_N / 32 - 1) {
          // the accumulator now lives in registers: hand the TMEM buffer back to the MMA thread
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tmem_empty_bar[buf]);
        }
        const int col0 = n0 + c * 32;
        if (row_ok) {
          bool done = false;
          if constexpr (OUT_BF16) {
            if (fast && col0 + 32 <= args.N) {
              done = true;
              uint4 o[4];
#pragma unroll
              for (int g = 0; g < 4; ++g) {
                float f[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) f[k] = __uint_as_float(v[g * 8 + k]);
                if (args.bias != nullptr) {
                  const float4 b0 = __ldg(reinterpret_cast<const float4*>(args.bias + col0 + g * 8));
                  const float4 b1 = __ldg(reinterpret_cast<const float4*>(args.bias + col0 + g * 8 + 4));
                  f[0] += b0.x, f[1] += b0.y, f[2] += b0.z, f[3] += b0.w;
                  f[4] += b1.x, f[5] += b1.y, f[6] += b1.z, f[7] += b1.w;
                }
                if (args.flags & kFlagRelu) {
#pragma unroll
                  for (int k = 0; k < 8; ++k) f[k] = fmaxf(f[k], 0.f);
                }
                if (add_row != nullptr) {
                  const uint32_t w[4] = {pa[g].x, pa[g].y, pa[g].z, pa[g].w};
#pragma unroll
                  for (int k = 0; k < 4; ++k)
                    f[2 * k] += __uint_as_float(w[k] << 16), f[2 * k + 1] += __uint_as_float(w[k] & 0xffff0000u);
                }
                if (mask_row != nullptr) {
                  const uint32_t w[4] = {pm[g].x, pm[g].y, pm[g].z, pm[g].w};
#pragma unroll
                  for (int k = 0; k < 4; ++k) {
                    f[2 * k] = __uint_as_float(w[k] << 16) > 0.f ? f[2 * k] : 0.f;
                    f[2 * k + 1] = __uint_as_float(w[k] & 0xffff0000u) > 0.f ? f[2 * k + 1] : 0.f;
                  }
                }
                o[g].x = pack_bf16x2(f[0], f[1]), o[g].y = pack_bf16x2(f[2], f[3]);
                o[g].z = pack_bf16x2(f[4], f[5]), o[g].w = pack_bf16x2(f[6], f[7]);
              }
              // operands of the next chunk: issued before this chunk's stores, in flight across the next TMEM load
              if (c + 1 < BLOCK_N / 32) prefetch(c + 1);
              uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(args.D) + (long long)row * args.ldd + col0);
#pragma unroll
              for (int g = 0; g < 4; ++g) dst[g] = o[g];
            } else if (wide_bf16 && col0 + 32 <= args.N) {
              done = true;
#pragma unroll
              for (int g = 0; g < 4; ++g) {
                float f[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) f[k] = __uint_as_float(v[g * 8 + k]);
                epilogue_store8_bf16(args, row, col0 + g * 8, f);
              }
            }
          }
          if (!done) {
#pragma unroll
            for (int g = 0; g < 8; ++g) {
              const int col = col0 + g * 4;
              if (col < args.N) {
                const float f0 = __uint_as_float(v[g * 4 + 0]), f1 = __uint_as_float(v[g * 4 + 1]);
                const float f2 = __uint_as_float(v[g * 4 + 2]), f3 = __uint_as_float(v[g * 4 + 3]);
                if (finalize)
                  red_add_f4(args.ws + (long long)row * args.ldws + col, f0, f1, f2, f3);
                else
                  epilogue_store4<OUT_BF16, SCATTER>(args, row, col, f0, f1, f2, f3);
              }
            }
          }
        }
      }
      if (finalize) {
        // Split-K rendezvous: the CTA that bumps the tile counter last owns the epilogue.
        const int et = threadIdx.x - 128;
        __threadfence();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (et == 0) {
          const uint32_t ctr = nt * args.m_tiles + mt;
          const uint32_t old = atomicAdd(&args.counters[ctr], 1u);
          *last_flag = (old == static_cast<uint32_t>(args.splits) - 1u) ? 1 : 0;
          if (*last_flag) args.counters[ctr] = 0;  // self-cleaning for the next launch
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (*last_flag) {
          __threadfence();
          // Only rows < M exist; batch the L2 reads (4 independent float4 per thread in flight)
          // instead of a load -> store -> load chain.
          constexpr int C4 = BLOCK_N / 4;
          const int rows = min(kBlockM, args.M - m0);
          const int total = rows * C4;
          for (int base = et; base < total; base += 128 * 4) {
            float4 v[4];
            int rr[4], cc[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int idx = base + u * 128;
              rr[u] = m0 + idx / C4, cc[u] = n0 + (idx % C4) * 4;
              if (idx < total && cc[u] < args.N)
                v[u] = __ldcg(reinterpret_cast<const float4*>(args.ws + (long long)rr[u] * args.ldws + cc[u]));
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int idx = base + u * 128;
              if (idx < total && cc[u] < args.N) {
                __stcg(reinterpret_cast<float4*>(args.ws + (long long)rr[u] * args.ldws + cc[u]),
                       make_float4(0.f, 0.f, 0.f, 0.f));
                epilogue_store4<OUT_BF16>(args, rr[u], cc[u], v[u].x, v[u].y, v[u].z, v[u].w);
              }
            }
          }
        }
      }
    }
  }

  if (dbg != nullptr && threadIdx.x == 128) dbg[6] = gtime_ns();     // epilogue done
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc_dyn(tmem_base, tmem_cols);
  if (dbg != nullptr && threadIdx.x == 0) dbg[7] = gtime_ns();
}


// ----------------------------------------------------------------------------------------
// CTA-pair GEMM (tcgen05 cta_group::2) for the long-K regime.
//
// A cluster of two CTAs (two SMs of one TPC) owns one 256 x BLOCK_N output tile.  CTA r of the pair loads
// rows [128 r, 128 r + 128) of the A tile and columns [BLOCK_N/2 r, ...) of the B tile into its own ring —
// half of B per SM, so each stage costs 16 KB + BLOCK_N/2 x 128 B instead of 16 KB + BLOCK_N x 128 B — and
// credits the bytes to the LEADER's full barrier.  One thread of the leader issues M = 256 MMAs that read both
// CTAs' shared memory; each CTA accumulates its 128 rows x BLOCK_N columns in its own tensor memory
// (double-buffered) and drains them with its own epilogue warps.  tcgen05.commit arrives on both CTAs'
// barriers.  No operand is multicast: every byte of A and B is loaded by exactly one CTA.
// ----------------------------------------------------------------------------------------
template <MmaKind KIND, bool A_MN, bool B_MN, int BLOCK_N, bool OUT_BF16>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(256, 1)
    gemm_tc_pair_kernel(const __grid_constant__ GemmArgs args) {
  constexpr int HALF_N = BLOCK_N / 2;
  using Cfg = GemmCfg<KIND, A_MN, B_MN, HALF_N>;           // per-CTA ring geometry
  constexpr uint32_t IDESC =
      umma_idesc(KIND == MmaKind::TF32 ? 2u : 1u, A_MN ? 1u : 0u, B_MN ? 1u : 0u, 2 * kBlockM, BLOCK_N);
  const int STAGES = args.stages;
  constexpr int STAGE_BYTES = Cfg::STAGE_BYTES;
  constexpr int BLOCK_K = Cfg::BLOCK_K;
  constexpr uint32_t ACC_COLS = BLOCK_N;                    // one accumulator: 128 lanes x BLOCK_N fp32 columns

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);   // used in the leader only
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;      // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;      // [2], used in the leader only (8 arrivals: 4 warps x 2 CTAs)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta_rank = cluster_ctarank();
  const bool leader = cta_rank == 0;
  const int pairs = gridDim.x >> 1, pair = blockIdx.x >> 1;
  const int total_tiles = args.m_tiles * args.n_tiles;       // m_tiles counts 256-row tiles here
  const uint32_t tmem_cols = total_tiles > pairs ? 2 * ACC_COLS : ACC_COLS;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&args.tmA);
    tma_prefetch_desc(&args.tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tmem_full_bar[b], 1);
      mbar_init(&tmem_empty_bar[b], 8);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc_pair(tmem_slot, tmem_cols);
  tc_fence_before();
  cluster_sync_all();          // both CTAs' barriers are initialised and both allocations are done
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();
  pdl_trigger();

  // Tile order: column groups of kGroupN tiles, rows fastest inside a group, so the tiles in flight at any
  // time (one per pair) share a few row blocks of A and a few column blocks of B that fit in L2 together;
  // B is then read from HBM once and A once per group instead of B once per wave.
  constexpr int kGroupN = 8;
  auto decode = [&](int tile, int& mt, int& nt) {
    const int per_group = kGroupN * args.m_tiles;
    const int grp = tile / per_group, r = tile - grp * per_group;
    const int width = min(kGroupN, args.n_tiles - grp * kGroupN);
    mt = r / width;
    nt = grp * kGroupN + (r - mt * width);
  };

  if (warp == 0) {
    // ===== TMA producer (both CTAs; bytes are credited to the leader's full barrier) =====
    if (lane == 0) {
      uint32_t it = 0;
      for (int tile = pair; tile < total_tiles; tile += pairs) {
        int mt, nt;
        decode(tile, mt, nt);
        const int m0 = mt * 2 * kBlockM + static_cast<int>(cta_rank) * kBlockM;
        const int n0 = nt * BLOCK_N + static_cast<int>(cta_rank) * HALF_N;
        for (int kb = 0; kb < args.kb_total; ++kb, ++it) {
          const int stage = it % STAGES;
          const uint32_t phase = (it / STAGES) & 1;
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * STAGE_BYTES;
          uint8_t* sb = sa + kABytes;
          const int k0 = kb * BLOCK_K;
          const uint32_t bar = mapa_u32(&full_bar[stage], 0);
          if (leader) mbar_expect_tx(&full_bar[stage], 2 * STAGE_BYTES);
          if constexpr (!A_MN) {
            tma_load_2d_pair(sa, &args.tmA, bar, k0, m0);
          } else {
#pragma unroll
            for (int a = 0; a < kBlockM / Cfg::ATOM_E; ++a)
              tma_load_2d_pair(sa + a * (BLOCK_K * 128), &args.tmA, bar, m0 + a * Cfg::ATOM_E, k0);
          }
          if constexpr (!B_MN) {
            tma_load_2d_pair(sb, &args.tmB, bar, k0, n0);
          } else {
#pragma unroll
            for (int a = 0; a < HALF_N / Cfg::ATOM_E; ++a)
              tma_load_2d_pair(sb + a * (BLOCK_K * 128), &args.tmB, bar, n0 + a * Cfg::ATOM_E, k0);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: one thread of the leader CTA drives both SMs' tensor cores =====
    if (leader && lane == 0) {
      uint32_t it = 0, t = 0;
      for (int tile = pair; tile < total_tiles; tile += pairs, ++t) {
        const uint32_t buf = t & 1;
        mbar_wait(&tmem_empty_bar[buf], ((t >> 1) & 1) ^ 1);   // both CTAs' epilogues have drained this accumulator
        tc_fence_after();
        const uint32_t acc = tmem_base + buf * ACC_COLS;
        for (int kb = 0; kb < args.kb_total; ++kb, ++it) {
          const int stage = it % STAGES;
          const uint32_t phase = (it / STAGES) & 1;
          mbar_wait(&full_bar[stage], phase);                    // both CTAs' halves have landed
          tc_fence_after();
          const uint32_t a_base = smem_u32(smem + stage * STAGE_BYTES);
          const uint32_t b_base = a_base + kABytes;
#pragma unroll
          for (int j = 0; j < BLOCK_K / Cfg::UMMA_K; ++j) {
            const uint64_t da = A_MN ? umma_smem_desc_sw128(a_base + j * Cfg::UMMA_K * 128,
                                                            BLOCK_K * 128, Cfg::MN_SBO, Cfg::MN_LAYOUT)
                                     : umma_smem_desc_sw128(a_base + j * 32, 16, 1024);
            const uint64_t db = B_MN ? umma_smem_desc_sw128(b_base + j * Cfg::UMMA_K * 128,
                                                            BLOCK_K * 128, Cfg::MN_SBO, Cfg::MN_LAYOUT)
                                     : umma_smem_desc_sw128(b_base + j * 32, 16, 1024);
            tc_mma_pair<KIND>(acc, da, db, IDESC, (kb | j) != 0 ? 1u : 0u);
          }
          tc_commit_pair(&empty_bar[stage], 0b11);   // frees this ring slot in BOTH CTAs
        }
        tc_commit_pair(&tmem_full_bar[buf], 0b11);   // accumulator complete in both CTAs
      }
    }
  } else if (warp >= 4) {
    // ===== epilogue: each CTA drains its own 128 rows; warp q owns TMEM lanes [32q, 32q+32) =====
    const int q = warp - 4;
    const bool wide_bf16 = OUT_BF16 && (args.flags & (kFlagOutBf16 | kFlagAccumulate)) == kFlagOutBf16 &&
                           (args.ldd & 7) == 0 && (args.mask == nullptr || (args.ldmask & 7) == 0);
    uint32_t t = 0;
    for (int tile = pair; tile < total_tiles; tile += pairs, ++t) {
      int mt, nt;
      decode(tile, mt, nt);
      const int n0 = nt * BLOCK_N;
      const int row = mt * 2 * kBlockM + static_cast<int>(cta_rank) * kBlockM + q * 32 + lane;
      const bool row_ok = row < args.M;
      const uint32_t buf = t & 1;
      mbar_wait(&tmem_full_bar[buf], (t >> 1) & 1);
      tc_fence_after();
      const uint32_t acc = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * ACC_COLS;
#pragma unroll 1
      for (int c = 0; c < BLOCK_N / 32; ++c) {
        uint32_t v[32];
        tmem_ld_32x32(acc + c * 32, v);
        tmem_ld_wait();
        if (c == BLOCK_N / 32 - 1) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(&tmem_empty_bar[buf], 0);   // the leader's barrier counts 8 warps
        }
        const int col0 = n0 + c * 32;
        if (!row_ok) continue;
        if (wide_bf16 && col0 + 32 <= args.N) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            float f[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) f[k] = __uint_as_float(v[g * 8 + k]);
            epilogue_store8_bf16(args, row, col0 + g * 8, f);
          }
        } else {
#pragma unroll
          for (int g = 0; g < 8; ++g) {
            const int col = col0 + g * 4;
            if (col < args.N)
              epilogue_store4<OUT_BF16>(args, row, col, __uint_as_float(v[g * 4 + 0]), __uint_as_float(v[g * 4 + 1]),
                                        __uint_as_float(v[g * 4 + 2]), __uint_as_float(v[g * 4 + 3]));
          }
        }
      }
    }
  }

  tc_fence_before();
  cluster_sync_all();          // the leader's MMAs read the peer's shared memory until the last commit
  if (warp == 2) tmem_dealloc_pair(tmem_base, tmem_cols);
}

// Debug hook: attach a timeline buffer (8 x int64 per CTA) to a plan, or detach with NULL.
// ---------------------------------------------------------------------------------------
// Host side
// ---------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int encode_tmap_2d(CUtensorMap* tm, int dtype, const void* base, uint64_t inner, uint64_t outer,
                   uint64_t row_stride_bytes, uint32_t box_inner, uint32_t box_outer,
                   bool mn_major) {
  static EncodeTiledFn fn = reinterpret_cast<EncodeTiledFn>(driver_fn("cuTensorMapEncodeTiled"));
  if (fn == nullptr) return B2D_ERR_UNSUPPORTED;
  const cuuint64_t dims[2] = {inner, outer};
  const cuuint64_t strides[1] = {row_stride_bytes};
  const cuuint32_t box[2] = {box_inner, box_outer};
  const cuuint32_t estr[2] = {1, 1};
  // fp32 operands are loaded as TFLOAT32: the TMA unit rounds them to tf32 on the way into
  // shared memory (round-to-nearest) instead of the MMA truncating the low mantissa bits.
  // B2D_TF32_TRUNCATE=1 keeps the raw fp32 bits (for A/B comparison of the rounding mode).
  static const bool truncate = getenv("B2D_TF32_TRUNCATE") != nullptr;
  const CUtensorMapDataType dt =
      dtype == B2D_F32 ? (truncate ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_TFLOAT32)
                       : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  // MN-major fp32 tiles are consumed with the 32 B-chunk swizzle (see GemmCfg::MN_LAYOUT).
  const CUtensorMapSwizzle sw = (mn_major && dtype == B2D_F32) ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B
                                                               : CU_TENSOR_MAP_SWIZZLE_128B;
  CUresult r = fn(tm, dt, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed: CUresult %d (inner=%llu outer=%llu stride=%llu box=%ux%u)",
              (int)r, (unsigned long long)inner, (unsigned long long)outer,
              (unsigned long long)row_stride_bytes, box_inner, box_outer);
    return B2D_ERR_CUDA;
  }
  return B2D_OK;
}

// Latency-bound regime (convolution-sized GEMMs: a handful of k-blocks per tile, several waves of
// tiles): what hides a CTA's serial prologue / epilogue is ANOTHER CTA on the same SM, not a deeper
// ring, so the ring is cut until 2-3 CTAs fit in shared memory together.  B2D_GEMM_CORESIDENT=0 keeps
// the one-CTA-per-SM deep-ring configuration everywhere (A/B switch).
static bool coresident_enabled() {
  static const bool on = [] {
    const char* e = getenv("B2D_GEMM_CORESIDENT");
    return e == nullptr || e[0] != '0';
  }();
  return on;
}
constexpr int kShallowKBlocks = 64;   // K loops up to this many k-blocks count as latency-bound

static int pick_block_n(const b2d_gemm_desc& d) {
  // A bf16 MN-major B tile is built from 64-element (128 B) atoms, so its minimum width is 64.
  if (d.N <= 32 && !(d.dtype == B2D_BF16 && d.b_mn_major)) return 32;
  if (d.N <= 64) return 64;
  const int mt = ceil_div(d.M, kBlockM);
  // Prefer the widest tile that still yields at least one full wave of CTAs (never wider than N).
  const int kb = ceil_div(d.K, d.dtype == B2D_F32 ? 32 : 64);
  const bool shallow = coresident_enabled() && kb <= kShallowKBlocks;
  if (d.N > 128 && mt * ceil_div(d.N, 256) >= sm_count() && !shallow) return 256;
  if (!shallow && d.N > 64 && d.splits <= 0 && !d.accumulate) {
    // Long K, too few tiles for a wave of wide ones: the main loop is bound by the bytes each CTA pulls into
    // shared memory per k-block (16 KB of A + 128 B per column of B), so prefer the tile width — with the
    // split-K factor that fills the SMs — that minimises waves x k-blocks per CTA x bytes per k-block.
    // (Dense-1024 forward at batch 1024: 128 x 64 tiles, 45.8 us -> 256-wide tiles, split-K 4.)
    const int sms = sm_count();
    int best_bn = 0;
    long long best_cost = 0;
    for (int bn : {256, 128, 64}) {
      if (bn > 64 && bn / 2 >= d.N) continue;
      const int tiles = mt * ceil_div(d.N, bn);
      const int splits = (tiles < sms / 2 && kb >= 8) ? max(1, min(kb / 2, sms / tiles)) : 1;
      const int per = ceil_div(kb, splits);
      const long long waves = ceil_div(tiles * ceil_div(kb, per), sms);
      const long long cost = waves * (per + (splits > 1 ? 6 : 2)) * (16 + bn / 8);   // + prologue / epilogue / meeting
      if (best_bn == 0 || cost < best_cost) best_bn = bn, best_cost = cost;
    }
    return best_bn;
  }
  if (mt * ceil_div(d.N, 128) >= sm_count() * 13 / 16) return 128;
  // Few tiles: narrower tiles double the CTA count and halve each CTA's epilogue.
  if (mt * ceil_div(d.N, 128) < sm_count() / 2) return 64;
  return d.N % 128 == 0 || d.N > 512 ? 128 : 64;
}

// Persistent CTAs to launch for a resolved geometry: one per tile up to what the GPU holds at once — two per
// SM when both their shared memory and their (double-buffered) accumulators fit together, else one.
static int persistent_ctas(const b2d_gemm_geometry& g) {
  const int total = g.m_tiles * g.n_tiles * g.splits;
  const int acc_cols = g.block_n < 32 ? 32 : g.block_n;
  const bool two = g.smem_bytes + 1024 <= (227 * 1024) / 2 && 2 * (2 * acc_cols) <= 512;
  return min(total, sm_count() * (two ? 2 : 1));
}

static int resolve(const b2d_gemm_desc& d, b2d_gemm_geometry* g) {
  B2D_REQUIRE(d.M > 0 && d.N > 0 && d.K > 0, "gemm: M, N, K must be positive");
  B2D_REQUIRE(d.dtype == B2D_F32 || d.dtype == B2D_BF16, "gemm: dtype must be F32 or BF16");
  B2D_REQUIRE(d.out_dtype == B2D_F32 || d.out_dtype == B2D_BF16, "gemm: bad out_dtype");
  const int es = d.dtype == B2D_F32 ? 4 : 2;
  B2D_REQUIRE((d.lda * es) % 16 == 0 && (d.ldb * es) % 16 == 0,
              "gemm: operand rows must be 16-byte aligned (lda=%lld ldb=%lld)", (long long)d.lda,
              (long long)d.ldb);
  B2D_REQUIRE(d.N % 4 == 0 && d.ldd % 4 == 0, "gemm: N and ldd must be multiples of 4");
  B2D_REQUIRE(!(d.accumulate && d.out_dtype != B2D_F32), "gemm: accumulate needs fp32 output");
  B2D_REQUIRE(d.mask == nullptr || d.ldmask % 4 == 0, "gemm: ldmask must be a multiple of 4");
  g->pair = 0;
  g->block_n = pick_block_n(d);
  g->m_tiles = ceil_div(d.M, kBlockM);
  g->n_tiles = ceil_div(d.N, g->block_n);
  const int block_k = 128 / es;
  g->k_blocks = ceil_div(d.K, block_k);
  int splits = d.splits;
  if (splits <= 0) {
    const int tiles = g->m_tiles * g->n_tiles;
    splits = 1;
    // Few tiles and a long K: spread K over the idle SMs, keeping >= 2 k-blocks per CTA
    // (B2D_SPLITK_MIN_KB overrides the 2: fewer, longer splits mean fewer partial tiles to meet).
    static const int min_kb = [] {
      const char* e = getenv("B2D_SPLITK_MIN_KB");
      const int v = e != nullptr ? atoi(e) : 2;
      return v >= 1 ? v : 2;
    }();
    if (tiles < sm_count() / 2 && g->k_blocks >= 8)
      splits = min(max(1, g->k_blocks / min_kb), max(1, sm_count() / tiles));
  }
  splits = max(1, min(splits, g->k_blocks));
  const int per = ceil_div(g->k_blocks, splits);
  g->splits = ceil_div(g->k_blocks, per);
  const int stage_bytes = kABytes + g->block_n * 128;
  // Ring depth: never deeper than the K loop of one CTA (short loops keep the kernel's
  // shared-memory footprint small), never deeper than fits.
  const int per_cta = ceil_div(g->k_blocks, g->splits);
  g->stages = max(2, min(min(kMaxStages, kSmemBudget / stage_bytes), per_cta));
  const int ctas = g->m_tiles * g->n_tiles * g->splits;
  if (coresident_enabled() && ctas > sm_count() && g->k_blocks <= kShallowKBlocks && g->block_n <= 128) {
    const int together = 2;   // registers (<= 128 per thread) allow two CTAs per SM
    const int budget = (227 * 1024) / together - 1024 /*per-CTA reserve*/ - 1024 - 256;
    g->stages = max(2, min(g->stages, budget / stage_bytes));
  }
  g->smem_bytes = g->stages * stage_bytes + 1024 + 256;
  g->ctas = persistent_ctas(*g);
  return B2D_OK;
}

// CTA-pair kernel (gemm_tc_pair_kernel): long-K plain GEMMs with at least one full wave of 256-row tiles.
// Picks the tile width (256 or 128 columns) that wastes the least of the last wave.  B2D_GEMM_PAIR=0 disables.
static void pick_pair(const b2d_gemm_desc& d, b2d_gemm_geometry* g) {
  static const bool on = [] {
    const char* e = getenv("B2D_GEMM_PAIR");
    return e == nullptr || e[0] != '0';
  }();
  g->pair = 0;
  if (!on || g->splits != 1 || g->k_blocks < 16 || d.M < 256 || d.N < 128) return;
  const int pairs = sm_count() / 2;
  const int mt2 = ceil_div(d.M, 2 * kBlockM);
  int best_bn = 0;
  double best_eff = 0.0;
  // Only 256-column tiles: with N = 128 per instruction each SM's tensor core re-reads its A rows twice as often
  // and falls far below peak (measured 808 vs 1083 TFLOP/s at 4096^3 bf16 against the single-CTA kernel).
  for (int bn : {256}) {
    if (bn > d.N) continue;
    const int tiles = mt2 * ceil_div(d.N, bn);
    if (tiles < pairs) continue;
    const double eff = static_cast<double>(tiles) / (static_cast<double>(ceil_div(tiles, pairs)) * pairs);
    if (eff >= 0.8 && eff > best_eff) best_eff = eff, best_bn = bn;
  }
  if (best_bn == 0) return;
  g->pair = 1;
  g->block_n = best_bn;
  g->m_tiles = mt2;
  g->n_tiles = ceil_div(d.N, best_bn);
  const int stage_bytes = kABytes + (best_bn / 2) * 128;
  g->stages = max(2, min(min(kMaxStages, kSmemBudget / stage_bytes), g->k_blocks));
  g->smem_bytes = g->stages * stage_bytes + 1024 + 256;
  g->ctas = 2 * min(g->m_tiles * g->n_tiles, pairs);
}

}  // namespace b2d

using namespace b2d;

struct b2d_gemm_plan {
  GemmArgs args;
  b2d_gemm_geometry geo;
  void (*kernel)(const GemmArgs);
  int smem_bytes;     // maximum the kernel instantiation may ask for (function attribute)
  int launch_smem;    // what this plan actually launches with (ring depth of this problem)
  int grid;           // persistent CTAs: min(tiles, SMs x CTAs that fit an SM together)
  int dtype, a_mn, b_mn, out_dtype;   // what the kernel was instantiated for (b2d_gemm_plan_set_scatter re-selects)
};

// Last step of every plan constructor: publish the tile grid to the kernel and size the launch.
static void finish_plan(b2d_gemm_plan* p) {
  GemmArgs& a = p->args;
  a.m_tiles = p->geo.m_tiles, a.n_tiles = p->geo.n_tiles;
  if (!p->geo.pair) p->geo.ctas = persistent_ctas(p->geo);      // the convolution plans changed m_tiles after resolve()
  p->grid = p->geo.ctas;
}

template <MmaKind KIND, bool A_MN, bool B_MN, bool OUT_BF16>
static void select_pair_n(int block_n, void (**k)(const GemmArgs), int* smem) {
  if (block_n == 256) {
    *k = gemm_tc_pair_kernel<KIND, A_MN, B_MN, 256, OUT_BF16>;
    *smem = GemmCfg<KIND, A_MN, B_MN, 128>::SMEM_BYTES;
  } else {
    *k = gemm_tc_pair_kernel<KIND, A_MN, B_MN, 128, OUT_BF16>;
    *smem = GemmCfg<KIND, A_MN, B_MN, 64>::SMEM_BYTES;
  }
}
template <MmaKind KIND>
static void select_pair(bool a_mn, bool b_mn, int block_n, int out_dtype, void (**k)(const GemmArgs), int* smem) {
  if (out_dtype == B2D_BF16) {
    if (a_mn && b_mn) return select_pair_n<KIND, true, true, true>(block_n, k, smem);
    if (a_mn) return select_pair_n<KIND, true, false, true>(block_n, k, smem);
    if (b_mn) return select_pair_n<KIND, false, true, true>(block_n, k, smem);
    return select_pair_n<KIND, false, false, true>(block_n, k, smem);
  }
  if (a_mn && b_mn) return select_pair_n<KIND, true, true, false>(block_n, k, smem);
  if (a_mn) return select_pair_n<KIND, true, false, false>(block_n, k, smem);
  if (b_mn) return select_pair_n<KIND, false, true, false>(block_n, k, smem);
  return select_pair_n<KIND, false, false, false>(block_n, k, smem);
}

template <MmaKind KIND, bool A_MN, bool B_MN, bool OUT_BF16>
static void select_n(int block_n, void (**k)(const GemmArgs), int* smem) {
  switch (block_n) {
    case 32:
      *k = gemm_tc_kernel<KIND, A_MN, B_MN, 32, OUT_BF16>;
      *smem = GemmCfg<KIND, A_MN, B_MN, 32>::SMEM_BYTES;
      break;
    case 64:
      *k = gemm_tc_kernel<KIND, A_MN, B_MN, 64, OUT_BF16>;
      *smem = GemmCfg<KIND, A_MN, B_MN, 64>::SMEM_BYTES;
      break;
    case 128:
      *k = gemm_tc_kernel<KIND, A_MN, B_MN, 128, OUT_BF16>;
      *smem = GemmCfg<KIND, A_MN, B_MN, 128>::SMEM_BYTES;
      break;
    default:
      *k = gemm_tc_kernel<KIND, A_MN, B_MN, 256, OUT_BF16>;
      *smem = GemmCfg<KIND, A_MN, B_MN, 256>::SMEM_BYTES;
      break;
  }
}

// The epilogue is compiled per output type: fp32-output kernels (every GEMM of the tf32 models) do not carry
// the bf16 store paths, which keeps their instruction footprint — what a run-once kernel pays for — small.
template <MmaKind KIND>
static void select_major(bool a_mn, bool b_mn, int block_n, int out_dtype, void (**k)(const GemmArgs), int* smem) {
  if (out_dtype == B2D_BF16) {
    if (a_mn && b_mn) return select_n<KIND, true, true, true>(block_n, k, smem);
    if (a_mn) return select_n<KIND, true, false, true>(block_n, k, smem);
    if (b_mn) return select_n<KIND, false, true, true>(block_n, k, smem);
    return select_n<KIND, false, false, true>(block_n, k, smem);
  }
  if (a_mn && b_mn) return select_n<KIND, true, true, false>(block_n, k, smem);
  if (a_mn) return select_n<KIND, true, false, false>(block_n, k, smem);
  if (b_mn) return select_n<KIND, false, true, false>(block_n, k, smem);
  return select_n<KIND, false, false, false>(block_n, k, smem);
}

extern "C" int b2d_gemm_query(const b2d_gemm_desc* desc, b2d_gemm_geometry* out) {
  B2D_REQUIRE(desc != nullptr && out != nullptr, "gemm_query: null argument");
  const int rc = resolve(*desc, out);
  if (rc == B2D_OK) pick_pair(*desc, out);
  return rc;
}

extern "C" size_t b2d_gemm_workspace_bytes(const b2d_gemm_desc* desc) {
  b2d_gemm_geometry g;
  if (desc == nullptr || resolve(*desc, &g) != B2D_OK) return 0;
  if (g.splits <= 1 || desc->accumulate) return 0;
  return static_cast<size_t>(g.m_tiles) * kBlockM * g.n_tiles * g.block_n * sizeof(float);
}

extern "C" size_t b2d_gemm_counter_count(const b2d_gemm_desc* desc) {
  b2d_gemm_geometry g;
  if (desc == nullptr || resolve(*desc, &g) != B2D_OK) return 0;
  if (g.splits <= 1 || desc->accumulate) return 0;
  return static_cast<size_t>(g.m_tiles) * g.n_tiles;
}

extern "C" int b2d_gemm_plan_create(const b2d_gemm_desc* desc, b2d_gemm_plan** out) {
  B2D_REQUIRE(desc != nullptr && out != nullptr, "gemm_plan_create: null argument");
  const b2d_gemm_desc& d = *desc;
  b2d_gemm_geometry g;
  int rc = resolve(d, &g);
  if (rc != B2D_OK) return rc;
  pick_pair(d, &g);
  const int b_box_n = g.pair ? g.block_n / 2 : g.block_n;      // a CTA of a pair loads half of the B tile
  B2D_REQUIRE(d.A && d.B && d.D, "gemm: null operand pointer");
  B2D_REQUIRE((reinterpret_cast<uintptr_t>(d.A) % 16) == 0 && (reinterpret_cast<uintptr_t>(d.B) % 16) == 0 &&
                  (reinterpret_cast<uintptr_t>(d.D) % 16) == 0,
              "gemm: operand pointers must be 16-byte aligned");
  const bool finalize = g.splits > 1 && !d.accumulate;
  B2D_REQUIRE(!finalize || (d.workspace != nullptr && d.counters != nullptr),
              "gemm: split-K without accumulate needs workspace and counters");

  b2d_gemm_plan* p = new (std::nothrow) b2d_gemm_plan();
  B2D_REQUIRE(p != nullptr, "gemm: out of host memory");
  const int es = d.dtype == B2D_F32 ? 4 : 2;
  const uint32_t block_k = 128 / es, atom = 128 / es;
  // K-major operand: dims {K, rows}; box {BLOCK_K, tile rows}.
  // MN-major operand: dims {rows(MN), K}; box {one 128 B atom of MN, BLOCK_K}.
  if (!d.a_mn_major)
    rc = encode_tmap_2d(&p->args.tmA, d.dtype, d.A, d.K, d.M, d.lda * es, block_k, kBlockM, false);
  else
    rc = encode_tmap_2d(&p->args.tmA, d.dtype, d.A, d.M, d.K, d.lda * es, atom, block_k, true);
  if (rc == B2D_OK) {
    if (!d.b_mn_major)
      rc = encode_tmap_2d(&p->args.tmB, d.dtype, d.B, d.K, d.N, d.ldb * es, block_k, b_box_n, false);
    else
      rc = encode_tmap_2d(&p->args.tmB, d.dtype, d.B, d.N, d.K, d.ldb * es, atom, block_k, true);
  }
  if (rc != B2D_OK) {
    delete p;
    return rc;
  }
  GemmArgs& a = p->args;
  a.D = d.D, a.bias = d.bias, a.mask = d.mask, a.ws = d.workspace, a.counters = d.counters;
  a.ldd = d.ldd, a.ldmask = d.ldmask, a.ldws = static_cast<long long>(g.n_tiles) * g.block_n;
  a.M = d.M, a.N = d.N, a.K = d.K;
  a.kb_total = g.k_blocks, a.kb_per_split = ceil_div(g.k_blocks, g.splits), a.splits = g.splits;
  a.flags = (d.relu ? kFlagRelu : 0) | (d.accumulate ? kFlagAccumulate : 0) |
            (finalize ? kFlagFinalize : 0) | (d.out_dtype == B2D_BF16 ? kFlagOutBf16 : 0) |
            (d.mask != nullptr && d.mask_dtype == B2D_BF16 ? kFlagMaskBf16 : 0);
  a.stages = g.stages;
  p->launch_smem = g.smem_bytes;
  p->geo = g;
  p->dtype = d.dtype, p->a_mn = d.a_mn_major, p->b_mn = d.b_mn_major, p->out_dtype = d.out_dtype;
  if (g.pair) {
    if (d.dtype == B2D_F32)
      select_pair<MmaKind::TF32>(d.a_mn_major, d.b_mn_major, g.block_n, d.out_dtype, &p->kernel, &p->smem_bytes);
    else
      select_pair<MmaKind::F16>(d.a_mn_major, d.b_mn_major, g.block_n, d.out_dtype, &p->kernel, &p->smem_bytes);
  } else if (d.dtype == B2D_F32) {
    select_major<MmaKind::TF32>(d.a_mn_major, d.b_mn_major, g.block_n, d.out_dtype, &p->kernel, &p->smem_bytes);
  } else {
    select_major<MmaKind::F16>(d.a_mn_major, d.b_mn_major, g.block_n, d.out_dtype, &p->kernel, &p->smem_bytes);
  }
  cudaError_t e = cudaFuncSetAttribute(reinterpret_cast<const void*>(p->kernel),
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, p->smem_bytes);
  if (e != cudaSuccess) {
    set_error("gemm: cudaFuncSetAttribute(smem=%d) failed: %s", p->smem_bytes, cudaGetErrorString(e));
    delete p;
    return B2D_ERR_CUDA;
  }
  finish_plan(p);
  *out = p;
  return B2D_OK;
}

extern "C" int b2d_gemm_plan_run(const b2d_gemm_plan* p, void* stream) {
  B2D_REQUIRE(p != nullptr, "gemm_plan_run: null plan");
  B2D_CUDA_CHECK(launch_pdl(p->kernel, dim3(p->grid), dim3(256), p->launch_smem, static_cast<cudaStream_t>(stream),
                            p->args));
  return B2D_OK;
}

extern "C" void b2d_gemm_plan_destroy(b2d_gemm_plan* p) { delete p; }

// Turns an accumulating fp32 plan into GEMM -> reduce-scatter (see ScatterArgs).
extern "C" int b2d_gemm_plan_set_scatter(b2d_gemm_plan* p, const b2d_gemm_scatter* sc) {
  B2D_REQUIRE(p != nullptr && sc != nullptr, "gemm_plan_set_scatter: null argument");
  GemmArgs& a = p->args;
  B2D_REQUIRE((a.flags & kFlagAccumulate) != 0 && (a.flags & (kFlagOutBf16 | kFlagFinalize)) == 0,
              "gemm_plan_set_scatter: the plan must accumulate into an fp32 output");
  B2D_REQUIRE(sc->world >= 1 && sc->world <= 8 && sc->per > 0 && sc->per % 4 == 0 && sc->d_off % 4 == 0 &&
                  sc->seg_begin % 4 == 0 && sc->d_off >= sc->seg_begin,
              "gemm_plan_set_scatter: bad shard geometry");
  for (int r = 0; r < sc->world; ++r) {
    B2D_REQUIRE(sc->base[r] != nullptr, "gemm_plan_set_scatter: null arena of replica %d", r);
    a.sc.base[r] = static_cast<float*>(sc->base[r]);
  }
  B2D_REQUIRE(a.sc.base[sc->rank] + sc->d_off == static_cast<float*>(a.D),
              "gemm_plan_set_scatter: D is not at d_off inside this replica's arena");
  a.sc.d_off = sc->d_off, a.sc.seg_begin = sc->seg_begin, a.sc.per = sc->per, a.sc.world = sc->world;
  a.flags |= kFlagScatter;
  // The reduce-scatter epilogue is its own instantiation (the weight-gradient layout: both operands MN-major,
  // fp32 output), so every other GEMM of a step stays free of its instructions.
  B2D_REQUIRE(p->a_mn && p->b_mn && p->out_dtype == B2D_F32 && !p->geo.pair && a.cv.enabled == 0,
              "gemm_plan_set_scatter: only plain weight-gradient GEMMs (A and B MN-major, fp32 output) can scatter");
#define B2D_SCATTER_PICK(KIND)                                                          \
  switch (p->geo.block_n) {                                                             \
    case 32: p->kernel = gemm_tc_kernel<KIND, true, true, 32, false, true>; break;     \
    case 64: p->kernel = gemm_tc_kernel<KIND, true, true, 64, false, true>; break;     \
    case 128: p->kernel = gemm_tc_kernel<KIND, true, true, 128, false, true>; break;   \
    default: p->kernel = gemm_tc_kernel<KIND, true, true, 256, false, true>; break;    \
  }
  if (p->dtype == B2D_F32) {
    B2D_SCATTER_PICK(MmaKind::TF32)
  } else {
    B2D_SCATTER_PICK(MmaKind::F16)
  }
#undef B2D_SCATTER_PICK
  B2D_CUDA_CHECK(cudaFuncSetAttribute(reinterpret_cast<const void*>(p->kernel),
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, p->smem_bytes));
  return B2D_OK;
}

// Latency analysis: 16 globaltimer stamps per CTA are written to `buf` (start, prologue done,
// upstream complete, first tiles landed, MMAs issued, accumulator ready, epilogue done, end).
extern "C" void b2d_gemm_plan_set_timeline(b2d_gemm_plan* p, long long* buf) {
  if (p != nullptr) p->args.dbg = buf;
}

// ---------------------------------------------------------------------------------------
// Implicit-GEMM convolution plans (stride 1, NHWC x HWIO) on the same kernel.       [W1, W3]
//   forward : y[n,oh,ow,:]  = act( sum_{r,s} x[n,oh+r-ph,ow+s-pw,:] . W[r,s,:,:] + b )
//             A = x through a 4-D TMA box per filter tap, B = W viewed [KH*KW*Ci][Co] (MN-major)
//   dgrad   : dx[n,ih,iw,:] = sum_{r,s} dy[n,ih+ph-r,iw+pw-s,:] . W[r,s,:,:]^T  (x mask)
//             A = dy through the mirrored taps, B = the same weight matrix read K-major
// Reference: Conv2D layers (README.md:293); the reference ships no kernels.
// ---------------------------------------------------------------------------------------
// b2d_conv_desc is declared in include/b200dist.h

static int encode_tmap_4d(CUtensorMap* tm, int dtype, const void* base, uint64_t C, uint64_t W,
                          uint64_t H, uint64_t N, uint32_t box_c, uint32_t box_w, uint32_t box_h,
                          uint32_t box_n, bool mn_major = false, uint32_t stride = 1) {
  static EncodeTiledFn fn = reinterpret_cast<EncodeTiledFn>(driver_fn("cuTensorMapEncodeTiled"));
  if (fn == nullptr) return B2D_ERR_UNSUPPORTED;
  const uint64_t es = dtype == B2D_F32 ? 4 : 2;
  const cuuint64_t dims[4] = {C, W, H, N};
  const cuuint64_t strides[3] = {C * es, W * C * es, H * W * C * es};
  // Strided convolutions: the box spans box_w*stride x box_h*stride input pixels and the TMA unit
  // keeps every stride-th one (traversal stride), so shared memory still receives box_w x box_h.
  const cuuint32_t box[4] = {box_c, box_w * stride, box_h * stride, box_n};
  const cuuint32_t estr[4] = {1, stride, stride, 1};
  const CUtensorMapDataType dt =
      dtype == B2D_F32 ? CU_TENSOR_MAP_DATA_TYPE_TFLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  // channels-as-MN (wgrad) fp32 tiles need the 32-byte-chunk swizzle, like the 2-D MN-major maps
  const CUtensorMapSwizzle sw = (mn_major && dtype == B2D_F32) ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B
                                                               : CU_TENSOR_MAP_SWIZZLE_128B;
  CUresult r = fn(tm, dt, 4, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled(4d) failed: CUresult %d (C=%llu W=%llu H=%llu N=%llu box=%u,%u,%u,%u)",
              (int)r, (unsigned long long)C, (unsigned long long)W, (unsigned long long)H,
              (unsigned long long)N, box_c, box_w, box_h, box_n);
    return B2D_ERR_CUDA;
  }
  return B2D_OK;
}

// 1 if the tensor-core path can take this convolution (otherwise use the direct kernels).
static inline int conv_stride(const b2d_conv_desc& d) { return d.stride > 1 ? d.stride : 1; }
static inline int conv_out_h(const b2d_conv_desc& d) {
  return d.OH > 0 ? d.OH : (d.H + 2 * d.pad_h - d.KH) / conv_stride(d) + 1;
}
static inline int conv_out_w(const b2d_conv_desc& d) {
  return d.OW > 0 ? d.OW : (d.W + 2 * d.pad_w - d.KW) / conv_stride(d) + 1;
}

extern "C" int b2d_conv_tc_supported(const b2d_conv_desc* d, int dgrad) {
  if (d == nullptr) return 0;
  const int es = d->dtype == B2D_F32 ? 4 : 2;
  const int OH = conv_out_h(*d), OW = conv_out_w(*d);
  if (conv_stride(*d) > 1 && dgrad) return 0;     // strided data gradient: direct kernels
  if (OW * conv_stride(*d) > 256) return 0;        // TMA box limit
  const int k_ch = dgrad ? d->Co : d->Ci;       // reduction channels
  const int n_ch = dgrad ? d->Ci : d->Co;       // output channels
  const int out_w = dgrad ? d->W : OW;
  if (OH <= 0 || OW <= 0) return 0;
  if ((k_ch * es) % 128 != 0) return 0;          // whole 128-byte channel chunks per tap
  if (n_ch % 4 != 0 || (n_ch * es) % 16 != 0) return 0;
  if (out_w > 128 || out_w < 1) return 0;        // one output row must fit the 128-row M tile
  return 1;
}

static int conv_plan_create(const b2d_conv_desc& d, bool dgrad, b2d_gemm_plan** out) {
  B2D_REQUIRE(b2d_conv_tc_supported(&d, dgrad), "conv: shape not supported by the tensor-core path");
  B2D_REQUIRE(d.x && d.w && d.y, "conv: null tensor pointer");
  const int es = d.dtype == B2D_F32 ? 4 : 2;
  const int block_k = 128 / es;
  const int OH = conv_out_h(d), OW = conv_out_w(d);
  const int stride = dgrad ? 1 : conv_stride(d);
  // geometry of the *output* of this GEMM and of the tensor the taps slide over
  const int out_h = dgrad ? d.H : OH, out_w = dgrad ? d.W : OW;
  const int in_h = dgrad ? OH : d.H, in_w = dgrad ? OW : d.W;
  const int k_ch = dgrad ? d.Co : d.Ci, n_ch = dgrad ? d.Ci : d.Co;
  const void* in_ptr = dgrad ? d.y : d.x;
  void* out_ptr = dgrad ? const_cast<void*>(d.x) : d.y;

  b2d_gemm_desc gd{};
  gd.dtype = d.dtype, gd.out_dtype = d.out_dtype;
  gd.M = d.N * out_h * out_w, gd.N = n_ch, gd.K = d.KH * d.KW * k_ch;
  gd.lda = k_ch, gd.ldb = dgrad ? d.Co : d.Co, gd.ldd = n_ch, gd.ldmask = n_ch;
  gd.b_mn_major = dgrad ? 0 : 1;
  gd.splits = 1;
  b2d_gemm_geometry g;
  int rc = resolve(gd, &g);
  if (rc != B2D_OK) return rc;

  b2d_gemm_plan* p = new (std::nothrow) b2d_gemm_plan();
  B2D_REQUIRE(p != nullptr, "conv: out of host memory");
  ConvGeom& cv = p->args.cv;
  cv.enabled = 1, cv.kw = d.KW, cv.c_chunks = k_ch / block_k;
  cv.wo_t = out_w;
  cv.ho_t = min(min(out_h, kBlockM / out_w), 256 / stride);
  cv.n_t = 1;
  if (cv.ho_t >= out_h) {
    cv.ho_t = out_h;
    cv.n_t = max(1, min(d.N, kBlockM / (out_w * out_h)));
  }
  cv.rows_used = cv.wo_t * cv.ho_t * cv.n_t;
  cv.OH = out_h, cv.OW = out_w, cv.N_img = d.N;
  cv.tiles_h = ceil_div(out_h, cv.ho_t);
  cv.tap_sign = dgrad ? -1 : 1;
  cv.w_base = dgrad ? d.pad_w : -d.pad_w;
  cv.h_base = dgrad ? d.pad_h : -d.pad_h;
  cv.b_rows_per_tap = d.Ci;
  cv.stride = stride;
  cv.tap_r0 = cv.tap_s0 = 0, cv.tap_step = 1, cv.kw_full = d.KW;
  cv.out_step = 1, cv.out_off_h = cv.out_off_w = 0, cv.OH_full = out_h, cv.OW_full = out_w;
  g.m_tiles = cv.n_t == 1 ? d.N * cv.tiles_h : ceil_div(d.N, cv.n_t);

  rc = encode_tmap_4d(&p->args.tmA, d.dtype, in_ptr, k_ch, in_w, in_h, d.N, block_k, cv.wo_t, cv.ho_t, cv.n_t, false,
                      stride);
  if (rc == B2D_OK) {
    const uint64_t w_rows = (uint64_t)d.KH * d.KW * d.Ci;
    if (!dgrad)   // MN-major B: dims {Co, rows}, box {one 128-byte atom of Co, BLOCK_K rows}
      rc = encode_tmap_2d(&p->args.tmB, d.dtype, d.w, d.Co, w_rows, (uint64_t)d.Co * es, block_k, block_k, true);
    else          // K-major B: dims {Co (K), rows (N = ci)}, box {BLOCK_K, BLOCK_N rows}
      rc = encode_tmap_2d(&p->args.tmB, d.dtype, d.w, d.Co, w_rows, (uint64_t)d.Co * es, block_k, g.block_n, false);
  }
  if (rc != B2D_OK) {
    delete p;
    return rc;
  }
  GemmArgs& a = p->args;
  a.D = out_ptr, a.bias = dgrad ? nullptr : d.bias, a.mask = d.mask, a.ws = nullptr, a.counters = nullptr;
  a.addend = dgrad ? d.addend : nullptr;
  a.ldd = n_ch, a.ldmask = n_ch, a.ldws = 0;
  a.M = gd.M, a.N = gd.N, a.K = gd.K;
  a.kb_total = g.k_blocks, a.kb_per_split = g.k_blocks, a.splits = 1;
  a.flags = (d.relu ? kFlagRelu : 0) | (d.out_dtype == B2D_BF16 ? kFlagOutBf16 : 0) |
            (d.mask != nullptr && d.mask_dtype == B2D_BF16 ? kFlagMaskBf16 : 0);
  a.stages = g.stages;
  p->launch_smem = g.smem_bytes;
  p->geo = g;
  if (d.dtype == B2D_F32)
    select_major<MmaKind::TF32>(false, !dgrad, g.block_n, d.out_dtype, &p->kernel, &p->smem_bytes);
  else
    select_major<MmaKind::F16>(false, !dgrad, g.block_n, d.out_dtype, &p->kernel, &p->smem_bytes);
  cudaError_t e = cudaFuncSetAttribute(reinterpret_cast<const void*>(p->kernel),
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, p->smem_bytes);
  if (e != cudaSuccess) {
    set_error("conv: cudaFuncSetAttribute(smem=%d) failed: %s", p->smem_bytes, cudaGetErrorString(e));
    delete p;
    return B2D_ERR_CUDA;
  }
  finish_plan(p);
  *out = p;
  return B2D_OK;
}

// Strided data gradient, one output-parity class (a, b) in [0, stride)^2 per plan:
//   dx[n, s*i+a, s*j+b, :] = sum over the taps r == (a+ph) mod s, c == (b+pw) mod s of
//                            dy[n, i + (a+ph-r)/s, j + (b+pw-c)/s, :] . W[r, c, :, :]^T
// i.e. a unit-stride convolution of dy with a sub-sampled filter, written to every s-th pixel.
// *out is set to NULL (rc = OK) for a class without taps: those pixels are zero — the caller
// clears dx first (b2d_memset_async).
extern "C" int b2d_conv_dgrad_class_plan_create(const b2d_conv_desc* dp, int a, int b, b2d_gemm_plan** out) {
  B2D_REQUIRE(dp != nullptr && out != nullptr, "conv_dgrad_class_plan_create: null argument");
  const b2d_conv_desc& d = *dp;
  *out = nullptr;
  const int st = conv_stride(d);
  B2D_REQUIRE(a >= 0 && a < st && b >= 0 && b < st, "conv_dgrad_class: class (%d,%d) out of range", a, b);
  const int es = d.dtype == B2D_F32 ? 4 : 2;
  const int block_k = 128 / es;
  B2D_REQUIRE((d.Co * es) % 128 == 0 && d.Ci % 4 == 0, "conv_dgrad_class: channel counts not supported");
  const int OH = conv_out_h(d), OW = conv_out_w(d);
  const int Hc = (d.H - a + st - 1) / st, Wc = (d.W - b + st - 1) / st;   // pixels of this class
  if (Hc <= 0 || Wc <= 0) return B2D_OK;
  const int r0 = (a + d.pad_h) % st, s0 = (b + d.pad_w) % st;
  const int ntr = r0 < d.KH ? (d.KH - r0 + st - 1) / st : 0;
  const int nts = s0 < d.KW ? (d.KW - s0 + st - 1) / st : 0;
  if (ntr == 0 || nts == 0) return B2D_OK;
  B2D_REQUIRE(Wc <= kBlockM, "conv_dgrad_class: class width %d exceeds the M tile", Wc);

  b2d_gemm_desc gd{};
  gd.dtype = d.dtype, gd.out_dtype = d.out_dtype;
  gd.M = d.N * Hc * Wc, gd.N = d.Ci, gd.K = ntr * nts * d.Co;
  gd.lda = d.Co, gd.ldb = d.Co, gd.ldd = d.Ci, gd.ldmask = d.Ci;
  gd.b_mn_major = 0, gd.splits = 1;
  b2d_gemm_geometry g;
  int rc = resolve(gd, &g);
  if (rc != B2D_OK) return rc;
  b2d_gemm_plan* p = new (std::nothrow) b2d_gemm_plan();
  B2D_REQUIRE(p != nullptr, "conv_dgrad_class: out of host memory");
  ConvGeom& cv = p->args.cv;
  cv.enabled = 1, cv.kw = nts, cv.c_chunks = d.Co / block_k;
  cv.wo_t = Wc;
  cv.ho_t = min(Hc, kBlockM / Wc);
  cv.n_t = 1;
  if (cv.ho_t >= Hc) {
    cv.ho_t = Hc;
    cv.n_t = max(1, min(d.N, kBlockM / (Wc * Hc)));
  }
  cv.rows_used = cv.wo_t * cv.ho_t * cv.n_t;
  cv.OH = Hc, cv.OW = Wc, cv.N_img = d.N;
  cv.tiles_h = ceil_div(Hc, cv.ho_t);
  cv.tap_sign = -1;
  cv.h_base = (a + d.pad_h - r0) / st, cv.w_base = (b + d.pad_w - s0) / st;
  cv.b_rows_per_tap = d.Ci, cv.stride = 1;
  cv.tap_r0 = r0, cv.tap_s0 = s0, cv.tap_step = st, cv.kw_full = d.KW;
  cv.out_step = st, cv.out_off_h = a, cv.out_off_w = b, cv.OH_full = d.H, cv.OW_full = d.W;
  g.m_tiles = cv.n_t == 1 ? d.N * cv.tiles_h : ceil_div(d.N, cv.n_t);
  rc = encode_tmap_4d(&p->args.tmA, d.dtype, d.y, d.Co, OW, OH, d.N, block_k, cv.wo_t, cv.ho_t, cv.n_t);
  if (rc == B2D_OK)
    rc = encode_tmap_2d(&p->args.tmB, d.dtype, d.w, d.Co, (uint64_t)d.KH * d.KW * d.Ci, (uint64_t)d.Co * es, block_k,
                        g.block_n, false);
  if (rc != B2D_OK) {
    delete p;
    return rc;
  }
  GemmArgs& ar = p->args;
  ar.D = const_cast<void*>(d.x), ar.bias = nullptr, ar.mask = d.mask, ar.ws = nullptr, ar.counters = nullptr;
  ar.ldd = d.Ci, ar.ldmask = d.Ci, ar.ldws = 0;
  ar.M = d.N * d.H * d.W, ar.N = gd.N, ar.K = gd.K;
  ar.kb_total = g.k_blocks, ar.kb_per_split = g.k_blocks, ar.splits = 1;
  ar.flags = (d.out_dtype == B2D_BF16 ? kFlagOutBf16 : 0) |
            (d.mask != nullptr && d.mask_dtype == B2D_BF16 ? kFlagMaskBf16 : 0);
  ar.stages = g.stages;
  p->launch_smem = g.smem_bytes;
  p->geo = g;
  if (d.dtype == B2D_F32)
    select_major<MmaKind::TF32>(false, false, g.block_n, d.out_dtype, &p->kernel, &p->smem_bytes);
  else
    select_major<MmaKind::F16>(false, false, g.block_n, d.out_dtype, &p->kernel, &p->smem_bytes);
  cudaError_t e = cudaFuncSetAttribute(reinterpret_cast<const void*>(p->kernel),
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, p->smem_bytes);
  if (e != cudaSuccess) {
    set_error("conv_dgrad_class: cudaFuncSetAttribute failed: %s", cudaGetErrorString(e));
    delete p;
    return B2D_ERR_CUDA;
  }
  finish_plan(p);
  *out = p;
  return B2D_OK;
}

extern "C" int b2d_memset_async(void* ptr, int value, size_t bytes, void* stream) {
  B2D_CUDA_CHECK(cudaMemsetAsync(ptr, value, bytes, static_cast<cudaStream_t>(stream)));
  return B2D_OK;
}

extern "C" int b2d_conv_fwd_plan_create(const b2d_conv_desc* d, b2d_gemm_plan** out) {
  B2D_REQUIRE(d != nullptr && out != nullptr, "conv_fwd_plan_create: null argument");
  return conv_plan_create(*d, false, out);
}
extern "C" int b2d_conv_dgrad_plan_create(const b2d_conv_desc* d, b2d_gemm_plan** out) {
  B2D_REQUIRE(d != nullptr && out != nullptr, "conv_dgrad_plan_create: null argument");
  return conv_plan_create(*d, true, out);
}

// ---------------------------------------------------------------------------------------
// Convolution filter gradient on tcgen05.                                             [W2]
//   dW[r,s,ci,co] += sum_{n,oh,ow} x[n,oh+r-ph,ow+s-pw,ci] * dy[n,oh,ow,co]
// Per filter tap this is a GEMM with the *pixels* as the reduction dimension:
//   D[co, ci] = sum_pix dy[pix, co] * x_tap[pix, ci]
// Both operands are MN-major (channels contiguous, pixels along K) and both come straight from
// the NHWC tensors through 4-D TMA boxes of (ho_t x OW) pixels — the x box shifted by the tap.
// One CTA = one (tap, co tile, ci tile) x a chunk of pixel boxes; the fp32 accumulator stays
// in TMEM across the whole chunk and is added to dW with coalesced reductions (the TMEM row is
// co, which is the contiguous dimension of HWIO).
// ---------------------------------------------------------------------------------------
namespace b2d {

struct WgradArgs {
  CUtensorMap tmDy;  // dims {Co, OW, OH, N}
  CUtensorMap tmX;   // dims {Ci, W,  H,  N}
  float* dw;
  int Ci, Co, KW, pad_h, pad_w, stride;
  int rows;          // pixels per k-block = OW * ho_t
  int ho_t, tiles_h; // box height, boxes per image
  int tiles_total;   // N * tiles_h
  int tiles_per_cta;
  int m_tiles, n_tiles;
  int stages;
};

template <MmaKind KIND, int BLOCK_N>
__global__ void __launch_bounds__(256, 1) conv_wgrad_tc_kernel(const __grid_constant__ WgradArgs args) {
  constexpr int ES = (KIND == MmaKind::TF32) ? 4 : 2;
  constexpr int ATOM = 128 / ES;          // channels per 128-byte atom
  constexpr int UMMA_K = 32 / ES;         // pixels per MMA
  constexpr int A_ATOMS = kBlockM / ATOM;
  constexpr int B_ATOMS = BLOCK_N / ATOM;
  constexpr uint32_t MN_LAYOUT = (KIND == MmaKind::TF32) ? kUmmaSw128Base32 : kUmmaSw128;
  constexpr uint32_t MN_SBO = (KIND == MmaKind::TF32) ? 512 : 1024;
  constexpr uint32_t IDESC = umma_idesc(KIND == MmaKind::TF32 ? 2u : 1u, 1u, 1u, kBlockM, BLOCK_N);
  constexpr uint32_t TMEM_COLS = BLOCK_N < 32 ? 32 : BLOCK_N;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  const int R = args.rows;
  const int atom_bytes = R * 128;
  const int stage_bytes = (A_ATOMS + B_ATOMS) * atom_bytes;
  const int STAGES = args.stages;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * stage_bytes);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int id = blockIdx.x;
  const int n_tile = id % args.n_tiles;
  id /= args.n_tiles;
  const int m_tile = id % args.m_tiles;
  const int tap = id / args.m_tiles;
  const int r = tap / args.KW, sx = tap - r * args.KW;
  const int co0 = m_tile * kBlockM, ci0 = n_tile * BLOCK_N;
  const int t_begin = blockIdx.y * args.tiles_per_cta;
  const int t_end = min(args.tiles_total, t_begin + args.tiles_per_cta);
  const int num_kb = t_end - t_begin;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&args.tmDy);
    tma_prefetch_desc(&args.tmX);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    mbar_init(tmem_full_bar, 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();
  pdl_trigger();
  if (num_kb <= 0) {  // nothing to reduce for this CTA (uneven split): contribute nothing
    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc<TMEM_COLS>(tmem_base);
    return;
  }

  if (warp == 0) {
    if (lane == 0) {
      for (int i = 0; i < num_kb; ++i) {
        const int stage = i % STAGES;
        const uint32_t phase = (i / STAGES) & 1;
        mbar_wait(&empty_bar[stage], phase ^ 1);
        mbar_expect_tx(&full_bar[stage], stage_bytes);
        const int t = t_begin + i;
        const int n = t / args.tiles_h, oh0 = (t - n * args.tiles_h) * args.ho_t;
        uint8_t* sa = smem + stage * stage_bytes;
        uint8_t* sb = sa + A_ATOMS * atom_bytes;
#pragma unroll
        for (int a = 0; a < A_ATOMS; ++a)
          tma_load_4d(sa + a * atom_bytes, &args.tmDy, &full_bar[stage], co0 + a * ATOM, 0, oh0, n);
#pragma unroll
        for (int a = 0; a < B_ATOMS; ++a)
          tma_load_4d(sb + a * atom_bytes, &args.tmX, &full_bar[stage], ci0 + a * ATOM, sx - args.pad_w,
                      oh0 * args.stride + r - args.pad_h, n);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const int mmas = R / UMMA_K;
      for (int i = 0; i < num_kb; ++i) {
        const int stage = i % STAGES;
        const uint32_t phase = (i / STAGES) & 1;
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t a_base = smem_u32(smem + stage * stage_bytes);
        const uint32_t b_base = a_base + A_ATOMS * atom_bytes;
        for (int j = 0; j < mmas; ++j) {
          const uint64_t da = umma_smem_desc_sw128(a_base + j * UMMA_K * 128, atom_bytes, MN_SBO, MN_LAYOUT);
          const uint64_t db = umma_smem_desc_sw128(b_base + j * UMMA_K * 128, atom_bytes, MN_SBO, MN_LAYOUT);
          tc_mma<KIND>(tmem_base, da, db, IDESC, (i | j) != 0 ? 1u : 0u);
        }
        tc_commit(&empty_bar[stage]);
      }
      tc_commit(tmem_full_bar);
    }
  } else if (warp >= 4) {
    const int q = warp - 4;
    const int co = co0 + q * 32 + lane;
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    float* base = args.dw + (size_t)tap * args.Ci * args.Co;
#pragma unroll 1
    for (int c = 0; c < BLOCK_N / 32; ++c) {
      uint32_t v[32];
      tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + c * 32, v);
      tmem_ld_wait();
      if (co < args.Co) {
#pragma unroll
        for (int k = 0; k < 32; ++k) {
          const int ci = ci0 + c * 32 + k;
          if (ci < args.Ci) atomicAdd(base + (size_t)ci * args.Co + co, __uint_as_float(v[k]));
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<TMEM_COLS>(tmem_base);
}

}  // namespace b2d

struct b2d_conv_wgrad_plan {
  WgradArgs args;
  void (*kernel)(const WgradArgs);
  dim3 grid;
  int smem_bytes;
};

static int gcd_int(int a, int b) { return b == 0 ? a : gcd_int(b, a % b); }

extern "C" int b2d_conv_wgrad_plan_create(const b2d_conv_desc* dp, b2d_conv_wgrad_plan** out) {
  B2D_REQUIRE(dp != nullptr && out != nullptr, "conv_wgrad_plan_create: null argument");
  const b2d_conv_desc& d = *dp;
  const int es = d.dtype == B2D_F32 ? 4 : 2;
  const int atom = 128 / es, umma_k = 32 / es;
  const int OH = conv_out_h(d), OW = conv_out_w(d);
  const int stride = conv_stride(d);
  B2D_REQUIRE(OH > 0 && OW > 0 && OW * stride <= 256, "conv_wgrad: unsupported output size %dx%d", OH, OW);
  B2D_REQUIRE((d.Ci * es) % 16 == 0 && (d.Co * es) % 16 == 0,
              "conv_wgrad: channel counts must make pixels 16-byte aligned (Ci=%d Co=%d)", d.Ci, d.Co);
  B2D_REQUIRE(d.x && d.w && d.y, "conv_wgrad: null tensor pointer");
  b2d_conv_wgrad_plan* p = new (std::nothrow) b2d_conv_wgrad_plan();
  B2D_REQUIRE(p != nullptr, "conv_wgrad: out of host memory");
  WgradArgs& a = p->args;
  // Box height: smallest h making OW*h a multiple of the MMA K, scaled up towards ~96 pixels.
  const int h_unit = umma_k / gcd_int(OW, umma_k);
  const int block_n = d.Ci <= 32 && atom <= 32 ? 32 : (d.Ci <= 64 ? 64 : 128);
  a.m_tiles = ceil_div(d.Co, kBlockM), a.n_tiles = ceil_div(d.Ci, block_n);
  const int a_atoms = kBlockM / atom, b_atoms = block_n / atom;
  // Pixel box per pipeline stage: ~96 pixels and a ring as deep as fits for one CTA per SM; ~64 pixels and a
  // ring cut to half the shared memory when two CTAs are meant to share an SM (the grid below always aims at
  // two CTAs per SM), provided that still leaves a ring of at least two stages.
  int ho_t = 0, stage_bytes = 0;
  for (int together = coresident_enabled() ? 1 : 0; together >= 0; --together) {
    const int max_rows = together ? 64 : 96;
    ho_t = h_unit;
    while (OW * (ho_t + h_unit) <= max_rows && (ho_t + h_unit) * stride <= 256 && ho_t < OH) ho_t += h_unit;
    stage_bytes = (a_atoms + b_atoms) * OW * ho_t * 128;
    a.stages = max(1, min(4, kSmemBudget / stage_bytes));
    if (together) a.stages = min(a.stages, (227 * 1024 / 2 - 1024 - 1024 - 256) / stage_bytes);
    if (a.stages >= 2) break;
  }
  a.stages = max(1, a.stages);
  B2D_REQUIRE(ho_t * stride <= 256, "conv_wgrad: pixel box too tall for a strided TMA box");
  a.ho_t = ho_t, a.rows = OW * ho_t;
  a.tiles_h = ceil_div(OH, ho_t);
  a.tiles_total = d.N * a.tiles_h;
  a.Ci = d.Ci, a.Co = d.Co, a.KW = d.KW, a.pad_h = d.pad_h, a.pad_w = d.pad_w, a.stride = stride;
  a.dw = static_cast<float*>(const_cast<void*>(d.w));
  if (stage_bytes > kSmemBudget) {
    set_error("conv_wgrad: pixel box of %d rows does not fit shared memory", a.rows);
    delete p;
    return B2D_ERR_UNSUPPORTED;
  }
  p->smem_bytes = a.stages * stage_bytes + 1024 + 256;
  const int fixed = d.KH * d.KW * a.m_tiles * a.n_tiles;
  int chunks = max(1, min(a.tiles_total, (sm_count() * 2 + fixed - 1) / fixed));
  a.tiles_per_cta = ceil_div(a.tiles_total, chunks);
  chunks = ceil_div(a.tiles_total, a.tiles_per_cta);
  p->grid = dim3(fixed, chunks);
  int rc = encode_tmap_4d(&a.tmDy, d.dtype, d.y, d.Co, OW, OH, d.N, atom, OW, ho_t, 1, true);
  if (rc == B2D_OK) rc = encode_tmap_4d(&a.tmX, d.dtype, d.x, d.Ci, d.W, d.H, d.N, atom, OW, ho_t, 1, true, stride);
  if (rc != B2D_OK) {
    delete p;
    return rc;
  }
  if (d.dtype == B2D_F32)
    p->kernel = block_n == 32 ? conv_wgrad_tc_kernel<MmaKind::TF32, 32>
                : block_n == 64 ? conv_wgrad_tc_kernel<MmaKind::TF32, 64>
                                : conv_wgrad_tc_kernel<MmaKind::TF32, 128>;
  else
    p->kernel = block_n == 64 ? conv_wgrad_tc_kernel<MmaKind::F16, 64> : conv_wgrad_tc_kernel<MmaKind::F16, 128>;
  cudaError_t e = cudaFuncSetAttribute(reinterpret_cast<const void*>(p->kernel),
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  if (e != cudaSuccess) {
    set_error("conv_wgrad: cudaFuncSetAttribute failed: %s", cudaGetErrorString(e));
    delete p;
    return B2D_ERR_CUDA;
  }
  *out = p;
  return B2D_OK;
}

extern "C" int b2d_conv_wgrad_plan_run(const b2d_conv_wgrad_plan* p, void* stream) {
  B2D_REQUIRE(p != nullptr, "conv_wgrad_plan_run: null plan");
  B2D_CUDA_CHECK(launch_pdl(p->kernel, p->grid, dim3(256), p->smem_bytes, static_cast<cudaStream_t>(stream), p->args));
  return B2D_OK;
}

extern "C" void b2d_conv_wgrad_plan_destroy(b2d_conv_wgrad_plan* p) { delete p; }

// ---------------------------------------------------------------------------------------
// b2d_gemm_noalloc — the same GEMM for callers that must not allocate device memory.
//
// Some users of this library cannot call cudaMalloc at the point where they need a GEMM:
// a host application that has handed the whole GPU to its own memory pool (so a driver
// allocation would fail, or would fragment a budget somebody else accounts for), a plug-in
// running inside another runtime's callback where driver allocations are forbidden, or a
// latency-critical loop that cannot afford the device-wide synchronisation an allocation
// implies while other streams are busy.  The regular entry points push the split-K scratch
// (fp32 partial sums + tile counters) onto the caller; this one keeps it inside the library, in
// a fixed array that is part of the module image.
//
// Consequences the caller has to know about:
//   * Every GPU that loads this library carries the array (1 MiB + 1 KiB) until the process
//     exits, whether or not the entry point is ever used.
//   * The scratch is sized at compile time.  A problem whose split-K rendezvous needs more than
//     kStaticWorkspaceFloats partial sums or kStaticCounterWords output tiles is rejected with
//     B2D_ERR_LIMIT (use b2d_gemm_plan_create with your own workspace instead).
//   * All calls on a device share the one array.  Each call clears the part it is about to use
//     (stream-ordered, before the kernel), so leftovers of an earlier call never leak in — but
//     two calls that are in flight on the GPU at the same time (different streams, or the same
//     call from two host threads) race on it and produce garbage.  Serialise them.
//   * The first call on each device looks up the array's device address.  Make that first call
//     with the intended device current, and before any stream capture is begun.
// Nothing in here calls cudaMalloc / cudaFree or any allocating routine, on the host or the
// device; the descriptor, tensor maps and launch arguments live on the caller's stack.
// ---------------------------------------------------------------------------------------
#include <mutex>

constexpr size_t kStaticCounterWords = 256;            // one counter per output tile
constexpr size_t kStaticWorkspaceFloats = 256 * 1024;  // 1 MiB of fp32 partial sums

// In host code this name is only a handle for cudaGetSymbolAddress — it is NOT a device
// pointer and must never be passed to a kernel directly.
static __device__ __align__(256) float b2d_gemm_static_scratch[kStaticCounterWords + kStaticWorkspaceFloats];

namespace {

constexpr int kMaxDevices = 64;
float* g_static_scratch_addr[kMaxDevices] = {};
std::mutex g_static_scratch_mu;

// Device address of the scratch array on the current device (resolved once per device).
int static_scratch_for_current_device(float** out) {
  int dev = -1;
  B2D_CUDA_CHECK(cudaGetDevice(&dev));
  B2D_REQUIRE(dev >= 0 && dev < kMaxDevices, "gemm_noalloc: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_static_scratch_mu);
  if (g_static_scratch_addr[dev] == nullptr) {
    void* p = nullptr;
    B2D_CUDA_CHECK(cudaGetSymbolAddress(&p, b2d_gemm_static_scratch));
    B2D_REQUIRE(p != nullptr, "gemm_noalloc: cudaGetSymbolAddress returned a null address");
    g_static_scratch_addr[dev] = static_cast<float*>(p);
  }
  *out = g_static_scratch_addr[dev];
  return B2D_OK;
}

}  // namespace

extern "C" size_t b2d_gemm_noalloc_max_workspace_floats(void) { return kStaticWorkspaceFloats; }
extern "C" size_t b2d_gemm_noalloc_max_tiles(void) { return kStaticCounterWords; }

extern "C" int b2d_gemm_noalloc(const b2d_gemm_desc* desc, void* stream) {
  B2D_REQUIRE(desc != nullptr, "gemm_noalloc: null descriptor");
  const b2d_gemm_desc& d = *desc;
  b2d_gemm_geometry g;
  int rc = resolve(d, &g);
  if (rc != B2D_OK) return rc;
  B2D_REQUIRE(d.A && d.B && d.D, "gemm_noalloc: null operand pointer");
  B2D_REQUIRE((reinterpret_cast<uintptr_t>(d.A) % 16) == 0 && (reinterpret_cast<uintptr_t>(d.B) % 16) == 0 &&
                  (reinterpret_cast<uintptr_t>(d.D) % 16) == 0,
              "gemm_noalloc: operand pointers must be 16-byte aligned");
  const bool finalize = g.splits > 1 && !d.accumulate;
  const size_t ws_floats = finalize ? static_cast<size_t>(g.m_tiles) * kBlockM * g.n_tiles * g.block_n : 0;
  const size_t tiles = finalize ? static_cast<size_t>(g.m_tiles) * g.n_tiles : 0;
  if (ws_floats > kStaticWorkspaceFloats || tiles > kStaticCounterWords) {
    set_error("gemm_noalloc: this problem's split-K scratch (%zu partial sums, %zu tiles) exceeds the "
              "built-in maximum (%zu partial sums, %zu tiles); use b2d_gemm_plan_create with a "
              "caller-provided workspace",
              ws_floats, tiles, kStaticWorkspaceFloats, kStaticCounterWords);
    return B2D_ERR_LIMIT;
  }
  float* scratch = nullptr;
  rc = static_scratch_for_current_device(&scratch);
  if (rc != B2D_OK) return rc;
  uint32_t* counters = reinterpret_cast<uint32_t*>(scratch);
  float* workspace = scratch + kStaticCounterWords;

  b2d_gemm_plan plan{};  // on the stack: no host allocation either
  const int es = d.dtype == B2D_F32 ? 4 : 2;
  const uint32_t block_k = 128 / es, atom = 128 / es;
  if (!d.a_mn_major)
    rc = encode_tmap_2d(&plan.args.tmA, d.dtype, d.A, d.K, d.M, d.lda * es, block_k, kBlockM, false);
  else
    rc = encode_tmap_2d(&plan.args.tmA, d.dtype, d.A, d.M, d.K, d.lda * es, atom, block_k, true);
  if (rc != B2D_OK) return rc;
  if (!d.b_mn_major)
    rc = encode_tmap_2d(&plan.args.tmB, d.dtype, d.B, d.K, d.N, d.ldb * es, block_k, g.block_n, false);
  else
    rc = encode_tmap_2d(&plan.args.tmB, d.dtype, d.B, d.N, d.K, d.ldb * es, atom, block_k, true);
  if (rc != B2D_OK) return rc;
  GemmArgs& a = plan.args;
  a.D = d.D, a.bias = d.bias, a.mask = d.mask, a.ws = workspace, a.counters = counters;
  a.ldd = d.ldd, a.ldmask = d.ldmask, a.ldws = static_cast<long long>(g.n_tiles) * g.block_n;
  a.M = d.M, a.N = d.N, a.K = d.K;
  a.kb_total = g.k_blocks, a.kb_per_split = ceil_div(g.k_blocks, g.splits), a.splits = g.splits;
  a.flags = (d.relu ? kFlagRelu : 0) | (d.accumulate ? kFlagAccumulate : 0) |
            (finalize ? kFlagFinalize : 0) | (d.out_dtype == B2D_BF16 ? kFlagOutBf16 : 0) |
            (d.mask != nullptr && d.mask_dtype == B2D_BF16 ? kFlagMaskBf16 : 0);
  a.stages = g.stages;
  if (d.dtype == B2D_F32)
    select_major<MmaKind::TF32>(d.a_mn_major, d.b_mn_major, g.block_n, d.out_dtype, &plan.kernel, &plan.smem_bytes);
  else
    select_major<MmaKind::F16>(d.a_mn_major, d.b_mn_major, g.block_n, d.out_dtype, &plan.kernel, &plan.smem_bytes);
  B2D_CUDA_CHECK(cudaFuncSetAttribute(reinterpret_cast<const void*>(plan.kernel),
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, plan.smem_bytes));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (finalize) {
    // The rendezvous needs zeroed counters and partial sums.  The array still holds whatever
    // the previous call (possibly a different shape, possibly aborted) left behind.
    B2D_CUDA_CHECK(cudaMemsetAsync(counters, 0, tiles * sizeof(uint32_t), st));
    B2D_CUDA_CHECK(cudaMemsetAsync(workspace, 0, ws_floats * sizeof(float), st));
  }
  plan.geo = g, plan.launch_smem = g.smem_bytes;
  finish_plan(&plan);
  B2D_CUDA_CHECK(launch_pdl(plan.kernel, dim3(plan.grid), dim3(256), g.smem_bytes, st, plan.args));
  return B2D_OK;
}
