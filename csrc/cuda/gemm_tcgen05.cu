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
};

struct GemmArgs {
  CUtensorMap tmA;
  CUtensorMap tmB;
  void* D;
  const float* bias;
  const float* mask;
  float* ws;
  uint32_t* counters;
  long long ldd;
  long long ldmask;
  long long ldws;
  int M, N, K;
  int kb_total, kb_per_split, splits;
  int flags;
};

__device__ __forceinline__ void red_add_f4(float* p, float a, float b, float c, float d) {
  asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b),
               "f"(c), "f"(d)
               : "memory");
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

// Applies bias / ReLU / mask to four consecutive columns of one row and stores them.
__device__ __forceinline__ void epilogue_store4(const GemmArgs& a, int row, int col, float v0,
                                                float v1, float v2, float v3) {
  if (a.bias != nullptr) {
    const float4 b = *reinterpret_cast<const float4*>(a.bias + col);
    v0 += b.x, v1 += b.y, v2 += b.z, v3 += b.w;
  }
  if (a.flags & kFlagRelu) {
    v0 = fmaxf(v0, 0.f), v1 = fmaxf(v1, 0.f), v2 = fmaxf(v2, 0.f), v3 = fmaxf(v3, 0.f);
  }
  if (a.mask != nullptr) {
    const float4 m = *reinterpret_cast<const float4*>(a.mask + (long long)row * a.ldmask + col);
    v0 = m.x > 0.f ? v0 : 0.f, v1 = m.y > 0.f ? v1 : 0.f;
    v2 = m.z > 0.f ? v2 : 0.f, v3 = m.w > 0.f ? v3 : 0.f;
  }
  if (a.flags & kFlagAccumulate) {
    red_add_f4(reinterpret_cast<float*>(a.D) + (long long)row * a.ldd + col, v0, v1, v2, v3);
  } else if (a.flags & kFlagOutBf16) {
    uint2 p;
    p.x = pack_bf16x2(v0, v1);
    p.y = pack_bf16x2(v2, v3);
    *reinterpret_cast<uint2*>(reinterpret_cast<uint16_t*>(a.D) + (long long)row * a.ldd + col) = p;
  } else {
    *reinterpret_cast<float4*>(reinterpret_cast<float*>(a.D) + (long long)row * a.ldd + col) =
        make_float4(v0, v1, v2, v3);
  }
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
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
  static constexpr uint32_t TMEM_COLS = BLOCK_N < 32 ? 32 : BLOCK_N;
  // MN-major tiles: 16-bit operands use the plain 128 B swizzle (8-row atoms, 1024 B);
  // 32-bit operands must use the 32 B-chunk variant (4-row atoms, 512 B).
  static constexpr uint32_t MN_LAYOUT = (KIND == MmaKind::TF32) ? kUmmaSw128Base32 : kUmmaSw128;
  static constexpr uint32_t MN_SBO = (KIND == MmaKind::TF32) ? 512 : 1024;
  static constexpr uint32_t IDESC =
      umma_idesc(KIND == MmaKind::TF32 ? 2u : 1u, A_MN ? 1u : 0u, B_MN ? 1u : 0u, kBlockM, BLOCK_N);
};

template <MmaKind KIND, bool A_MN, bool B_MN, int BLOCK_N>
__global__ void __launch_bounds__(256, 1) gemm_tc_kernel(const __grid_constant__ GemmArgs args) {
  using Cfg = GemmCfg<KIND, A_MN, B_MN, BLOCK_N>;
  constexpr int STAGES = Cfg::STAGES;
  constexpr int STAGE_BYTES = Cfg::STAGE_BYTES;
  constexpr int BLOCK_K = Cfg::BLOCK_K;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);
  int* last_flag = reinterpret_cast<int*>(tmem_slot + 1);

  const int warp = threadIdx.x >> 5;  // warp-uniform
  const int lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * kBlockM;
  const int n0 = blockIdx.y * BLOCK_N;
  const int kb_begin = blockIdx.z * args.kb_per_split;
  const int kb_end = min(args.kb_total, kb_begin + args.kb_per_split);
  const int num_kb = kb_end - kb_begin;  // >= 1 by construction on the host

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&args.tmA);
    tma_prefetch_desc(&args.tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    mbar_init(tmem_full_bar, 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      for (int i = 0; i < num_kb; ++i) {
        const int stage = i % STAGES;
        const uint32_t phase = (i / STAGES) & 1;
        mbar_wait(&empty_bar[stage], phase ^ 1);
        mbar_expect_tx(&full_bar[stage], STAGE_BYTES);
        uint8_t* sa = smem + stage * STAGE_BYTES;
        uint8_t* sb = sa + kABytes;
        const int k0 = (kb_begin + i) * BLOCK_K;
        if constexpr (!A_MN) {
          tma_load_2d(sa, &args.tmA, &full_bar[stage], k0, m0);
        } else {
#pragma unroll
          for (int a = 0; a < kBlockM / Cfg::ATOM_E; ++a)
            tma_load_2d(sa + a * (BLOCK_K * 128), &args.tmA, &full_bar[stage], m0 + a * Cfg::ATOM_E,
                        k0);
        }
        if constexpr (!B_MN) {
          tma_load_2d(sb, &args.tmB, &full_bar[stage], k0, n0);
        } else {
#pragma unroll
          for (int a = 0; a < (BLOCK_N + Cfg::ATOM_E - 1) / Cfg::ATOM_E; ++a)
            tma_load_2d(sb + a * (BLOCK_K * 128), &args.tmB, &full_bar[stage], n0 + a * Cfg::ATOM_E,
                        k0);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (single thread) =====
    if (lane == 0) {
      for (int i = 0; i < num_kb; ++i) {
        const int stage = i % STAGES;
        const uint32_t phase = (i / STAGES) & 1;
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
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
          tc_mma<KIND>(tmem_base, da, db, Cfg::IDESC, (i | j) != 0 ? 1u : 0u);
        }
        tc_commit(&empty_bar[stage]);  // frees this ring slot once the MMAs above retire
      }
      tc_commit(tmem_full_bar);  // accumulator complete
    }
  } else if (warp >= 4) {
    // ===== epilogue: 4 warps, warp q owns TMEM lanes [32q, 32q+32) =====
    const int q = warp - 4;
    const int row = m0 + q * 32 + lane;
    const bool row_ok = row < args.M;
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    const bool finalize = (args.flags & kFlagFinalize) != 0;
#pragma unroll 1
    for (int c = 0; c < BLOCK_N / 32; ++c) {
      uint32_t v[32];
      tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + c * 32, v);
      tmem_ld_wait();
      const int col0 = n0 + c * 32;
      if (row_ok) {
#pragma unroll
        for (int g = 0; g < 8; ++g) {
          const int col = col0 + g * 4;
          if (col < args.N) {
            const float f0 = __uint_as_float(v[g * 4 + 0]), f1 = __uint_as_float(v[g * 4 + 1]);
            const float f2 = __uint_as_float(v[g * 4 + 2]), f3 = __uint_as_float(v[g * 4 + 3]);
            if (finalize)
              red_add_f4(args.ws + (long long)row * args.ldws + col, f0, f1, f2, f3);
            else
              epilogue_store4(args, row, col, f0, f1, f2, f3);
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
        const uint32_t tile = blockIdx.y * gridDim.x + blockIdx.x;
        const uint32_t old = atomicAdd(&args.counters[tile], 1u);
        *last_flag = (old == static_cast<uint32_t>(args.splits) - 1u) ? 1 : 0;
        if (*last_flag) args.counters[tile] = 0;  // self-cleaning for the next launch
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (*last_flag) {
        __threadfence();
        constexpr int C4 = BLOCK_N / 4;
        for (int idx = et; idx < kBlockM * C4; idx += 128) {
          const int r = m0 + idx / C4;
          const int col = n0 + (idx % C4) * 4;
          if (r < args.M && col < args.N) {
            float4* p = reinterpret_cast<float4*>(args.ws + (long long)r * args.ldws + col);
            const float4 s = __ldcg(p);
            __stcg(p, make_float4(0.f, 0.f, 0.f, 0.f));
            epilogue_store4(args, r, col, s.x, s.y, s.z, s.w);
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
}

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

static int pick_block_n(const b2d_gemm_desc& d) {
  // A bf16 MN-major B tile is built from 64-element (128 B) atoms, so its minimum width is 64.
  if (d.N <= 32 && !(d.dtype == B2D_BF16 && d.b_mn_major)) return 32;
  if (d.N <= 64) return 64;
  const int mt = ceil_div(d.M, kBlockM);
  // Prefer the widest tile that still yields at least one full wave of CTAs.
  if (mt * ceil_div(d.N, 256) >= 148) return 256;
  if (mt * ceil_div(d.N, 128) >= 120) return 128;
  return d.N % 128 == 0 || d.N > 512 ? 128 : 64;
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
  g->block_n = pick_block_n(d);
  g->m_tiles = ceil_div(d.M, kBlockM);
  g->n_tiles = ceil_div(d.N, g->block_n);
  const int block_k = 128 / es;
  g->k_blocks = ceil_div(d.K, block_k);
  int splits = d.splits;
  if (splits <= 0) {
    const int tiles = g->m_tiles * g->n_tiles;
    splits = 1;
    // Few tiles and a long K: spread K over the idle SMs, keeping >= 2 k-blocks per CTA.
    if (tiles < 74 && g->k_blocks >= 8) splits = min(g->k_blocks / 2, max(1, 148 / tiles));
  }
  splits = max(1, min(splits, g->k_blocks));
  const int per = ceil_div(g->k_blocks, splits);
  g->splits = ceil_div(g->k_blocks, per);
  const int stage_bytes = kABytes + g->block_n * 128;
  g->stages = min(kMaxStages, kSmemBudget / stage_bytes);
  g->smem_bytes = g->stages * stage_bytes + 1024 + 256;
  return B2D_OK;
}

}  // namespace b2d

using namespace b2d;

struct b2d_gemm_plan {
  GemmArgs args;
  b2d_gemm_geometry geo;
  void (*kernel)(const GemmArgs);
  int smem_bytes;
};

template <MmaKind KIND, bool A_MN, bool B_MN>
static void select_n(int block_n, void (**k)(const GemmArgs), int* smem) {
  switch (block_n) {
    case 32:
      *k = gemm_tc_kernel<KIND, A_MN, B_MN, 32>;
      *smem = GemmCfg<KIND, A_MN, B_MN, 32>::SMEM_BYTES;
      break;
    case 64:
      *k = gemm_tc_kernel<KIND, A_MN, B_MN, 64>;
      *smem = GemmCfg<KIND, A_MN, B_MN, 64>::SMEM_BYTES;
      break;
    case 128:
      *k = gemm_tc_kernel<KIND, A_MN, B_MN, 128>;
      *smem = GemmCfg<KIND, A_MN, B_MN, 128>::SMEM_BYTES;
      break;
    default:
      *k = gemm_tc_kernel<KIND, A_MN, B_MN, 256>;
      *smem = GemmCfg<KIND, A_MN, B_MN, 256>::SMEM_BYTES;
      break;
  }
}

template <MmaKind KIND>
static void select_major(bool a_mn, bool b_mn, int block_n, void (**k)(const GemmArgs), int* smem) {
  if (a_mn && b_mn) return select_n<KIND, true, true>(block_n, k, smem);
  if (a_mn) return select_n<KIND, true, false>(block_n, k, smem);
  if (b_mn) return select_n<KIND, false, true>(block_n, k, smem);
  return select_n<KIND, false, false>(block_n, k, smem);
}

extern "C" int b2d_gemm_query(const b2d_gemm_desc* desc, b2d_gemm_geometry* out) {
  B2D_REQUIRE(desc != nullptr && out != nullptr, "gemm_query: null argument");
  return resolve(*desc, out);
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
      rc = encode_tmap_2d(&p->args.tmB, d.dtype, d.B, d.K, d.N, d.ldb * es, block_k, g.block_n, false);
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
            (finalize ? kFlagFinalize : 0) | (d.out_dtype == B2D_BF16 ? kFlagOutBf16 : 0);
  p->geo = g;
  if (d.dtype == B2D_F32)
    select_major<MmaKind::TF32>(d.a_mn_major, d.b_mn_major, g.block_n, &p->kernel, &p->smem_bytes);
  else
    select_major<MmaKind::F16>(d.a_mn_major, d.b_mn_major, g.block_n, &p->kernel, &p->smem_bytes);
  cudaError_t e = cudaFuncSetAttribute(reinterpret_cast<const void*>(p->kernel),
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, p->smem_bytes);
  if (e != cudaSuccess) {
    set_error("gemm: cudaFuncSetAttribute(smem=%d) failed: %s", p->smem_bytes, cudaGetErrorString(e));
    delete p;
    return B2D_ERR_CUDA;
  }
  *out = p;
  return B2D_OK;
}

extern "C" int b2d_gemm_plan_run(const b2d_gemm_plan* p, void* stream) {
  B2D_REQUIRE(p != nullptr, "gemm_plan_run: null plan");
  const dim3 grid(p->geo.m_tiles, p->geo.n_tiles, p->geo.splits);
  p->kernel<<<grid, 256, p->smem_bytes, static_cast<cudaStream_t>(stream)>>>(p->args);
  count_launch();
  B2D_CUDA_CHECK(cudaGetLastError());
  return B2D_OK;
}

extern "C" void b2d_gemm_plan_destroy(b2d_gemm_plan* p) { delete p; }

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

  b2d_gemm_plan plan;  // on the stack: no host allocation either
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
            (finalize ? kFlagFinalize : 0) | (d.out_dtype == B2D_BF16 ? kFlagOutBf16 : 0);
  if (d.dtype == B2D_F32)
    select_major<MmaKind::TF32>(d.a_mn_major, d.b_mn_major, g.block_n, &plan.kernel, &plan.smem_bytes);
  else
    select_major<MmaKind::F16>(d.a_mn_major, d.b_mn_major, g.block_n, &plan.kernel, &plan.smem_bytes);
  B2D_CUDA_CHECK(cudaFuncSetAttribute(reinterpret_cast<const void*>(plan.kernel),
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, plan.smem_bytes));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (finalize) {
    // The rendezvous needs zeroed counters and partial sums.  The array still holds whatever
    // the previous call (possibly a different shape, possibly aborted) left behind.
    B2D_CUDA_CHECK(cudaMemsetAsync(counters, 0, tiles * sizeof(uint32_t), st));
    B2D_CUDA_CHECK(cudaMemsetAsync(workspace, 0, ws_floats * sizeof(float), st));
  }
  const dim3 grid(g.m_tiles, g.n_tiles, g.splits);
  plan.kernel<<<grid, 256, plan.smem_bytes, st>>>(plan.args);
  count_launch();
  B2D_CUDA_CHECK(cudaGetLastError());
  return B2D_OK;
}
