// allreduce.cu — the communication-bound hot path: cross-replica gradient all-reduce FUSED with
// the 1/num_workers averaging and the SGD update, over NVLink peer memory (sm_100a).
//
// Every rank holds a weight arena W_r and a gradient arena G_r in symmetric memory: every GPU
// can load/store every peer's arenas directly through NVSwitch (csrc/runtime/symm.cpp).
// One kernel per step replaces "batch_all_reduce: 6 all-reduces" + 4 metric all-reduces + the
// optimizer (README.md:403-407, 302).  Three algorithms, chosen per message size:
//
//   one-shot   every rank reads all peers' G, sums in rank order (bit-identical everywhere),
//              updates its own W.                          traffic/GPU: (N-1) S in
//   two-shot   rank r owns shard r: reads shard r of every G, averages, applies SGD, and writes
//              the *updated weights* into every peer's W — the all-gather carries weights, so
//              no optimizer pass exists at all.            traffic/GPU: (N-1)/N S in + out
//   pull       latency-optimised two-shot: rank r reduces shard r and updates its OWN weights only;
//              after a cheap flag (local fence, relaxed store) every rank *pulls* the other
//              shards of the updated weights.  No remote writes -> no system-scope release.
//   NVLS       like two-shot, but the switch does the sum (multimem.ld_reduce) and the
//              broadcast (multimem.st).                     traffic/GPU: S/N in + S/N out
//
// Elements >= n_params are the metric tail slots (loss sum, correct count, ...): they are
// summed, not averaged, and *added* to the running sums kept in W's tail.
// After the update each rank zeroes its own G, so the next backward can accumulate into it.
//
// Cross-GPU barriers are per block: block b of every rank meets block b of every other rank
// on monotone epoch counters in the peers' signal pads (st.release.sys / ld.acquire.sys).
#include "host_common.h"
#include "ptx.cuh"

namespace b2d {

constexpr int kMaxRanks = 8;
constexpr int kArThreads = 512;
constexpr int kMaxArBlocks = 144;  // signal pad: kMaxArBlocks * kMaxRanks words per rank

struct ArArgs {
  float* w[kMaxRanks];        // weight arenas (peer-mapped); w[rank] is local
  float* g[kMaxRanks];        // gradient arenas
  uint32_t* sig[kMaxRanks];   // signal pads
  uint32_t* epoch[kMaxRanks]; // per-rank local barrier epochs (only [rank] is used; all set in emulation)
  float* w_mc;                // NVLS multicast views (null when not bound)
  float* g_mc;
  float* out[kMaxRanks];      // plain all-reduce (no SGD): destination buffers
  long long n_params4;        // float4 count of the SGD range
  long long n_total4;         // float4 count of the whole arena
  float scale;                // lr / world
  int rank, world;
  int emulate;                // 1: blockIdx.y is the rank (single-GPU cooperative test mode)
  int plain;                  // 1: out = sum(g), no SGD, no zeroing
  long long seg_begin4[2];    // up to two [begin, end) float4 ranges of the arenas handled by this
  long long seg_end4[2];      // launch (bucketing); default one range = the whole arena
  int nseg;
  long long* dbg;             // optional timeline (globaltimer ns), 8 slots per block
  float* gz;                  // double-buffered gradients: the OTHER local buffer, zeroed during
                              // the reduction (null: single buffer, zeroed after a second barrier)
};

__device__ __forceinline__ long long ar_time_ns() {
  long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// Block-level barrier across ranks.  All threads of the block call it.
// `publish` = true when this block has written REMOTE memory since the last barrier (two-shot /
// NVLS weight scatter): the flag store must then be a system-scope release, which costs ~3 us.
// At the entry barrier nothing has been written by this kernel — the gradients the flag
// announces were produced by earlier kernels and are already complete in this GPU's L2, which is
// where peers read them — so a relaxed flag store and relaxed polling suffice (measured: 4.0 us
// -> see profiles/).
template <bool kPublish>
__device__ __forceinline__ void rank_barrier(const ArArgs& a, int rank, uint32_t* s_epoch,
                                             long long* dbg = nullptr) {
  __syncthreads();  // every thread's earlier accesses are ordered before the release below
  const int b = blockIdx.x;
  if (threadIdx.x == 0) {
    uint32_t* ep = a.epoch[rank] + b;
    *s_epoch = *ep + 1;
    *ep = *s_epoch;
  }
  __syncthreads();
  const uint32_t e = *s_epoch;
  if (threadIdx.x < a.world) {
    const int p = threadIdx.x;
    if (kPublish)
      st_release_sys(a.sig[p] + b * kMaxRanks + rank, e);
    else
      st_relaxed_sys(a.sig[p] + b * kMaxRanks + rank, e);
    if (dbg != nullptr && threadIdx.x == 0) dbg[0] = ar_time_ns();   // signals sent (dbg = &local[2])
    const uint32_t* mine = a.sig[rank] + b * kMaxRanks + p;
    long long spins = 0;
    while (static_cast<int32_t>((kPublish ? ld_acquire_sys(mine) : ld_relaxed_sys(mine)) - e) < 0) {
      if (++spins > (1ll << 31)) {  // a peer died: fail loudly instead of hanging the GPU forever
        printf("b200dist: rank %d block %d timed out waiting for rank %d (epoch %u)\n", rank, b, p, e);
        __trap();
      }
    }
  }
  __syncthreads();
}

__device__ __forceinline__ float4 f4_add(float4 a, float4 b) {
  return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
}

__device__ __forceinline__ float4 apply_update(float4 w, float4 s, bool is_param, float scale) {
  if (is_param) return make_float4(w.x - scale * s.x, w.y - scale * s.y, w.z - scale * s.z, w.w - scale * s.w);
  return f4_add(w, s);  // metric tail: running sums
}

enum ArAlgo : int { kOneShot = 0, kTwoShot = 1, kNvls = 2, kPull = 3 };

template <int ALGO>
__global__ void __launch_bounds__(kArThreads) allreduce_sgd_kernel(const ArArgs a) {
  __shared__ uint32_t s_epoch;
  const int rank = a.emulate ? blockIdx.y : a.rank;
  const int world = a.world;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);

  // Debug timeline: per block, slot 0 counts launches and slots 1..4 accumulate nanoseconds spent
  // (1) launched-early-to-upstream-complete, (2) in the release, (3) waiting for peers, (4) moving data.
  long long* dbg = a.dbg == nullptr ? nullptr : a.dbg + 8 * blockIdx.x;
  __shared__ long long s_t[4];
  if (dbg != nullptr && threadIdx.x == 0) s_t[0] = ar_time_ns();
  pdl_wait();     // this replica's backward has finished writing its gradients
  pdl_trigger();
  if (dbg != nullptr && threadIdx.x == 0) s_t[1] = ar_time_ns();
  rank_barrier<false>(a, rank, &s_epoch, dbg == nullptr ? nullptr : &s_t[2]);  // A: every replica's gradients are complete
  if (dbg != nullptr && threadIdx.x == 0) s_t[3] = ar_time_ns();

  // Peer loads have ~2 us latency.  All loads of a batch — every peer x kUnroll float4 — are issued
  // before the first use, and only then summed in fixed rank order (identical bits everywhere).
  constexpr int kUnroll = 2;
  auto gather_sum = [&](long long i0, long long limit, float4 (&acc)[kUnroll]) {
    float4 v[kMaxRanks][kUnroll];
#pragma unroll
    for (int p = 0; p < kMaxRanks; ++p)
      if (p < world) {
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
          const long long i = i0 + u * stride;
          v[p][u] = i < limit ? ld_nc_na_f4(reinterpret_cast<const float4*>(a.g[p]) + i) : zero;
        }
      }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) acc[u] = zero;
#pragma unroll
    for (int p = 0; p < kMaxRanks; ++p)
      if (p < world) {
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) acc[u] = f4_add(acc[u], v[p][u]);
      }
  };
#pragma unroll 1
  for (int sg = 0; sg < a.nseg; ++sg) {
  const long long seg_b = a.seg_begin4[sg], seg_e = a.seg_end4[sg];
  if (ALGO == kOneShot) {
    float4* wl = reinterpret_cast<float4*>(a.plain ? a.out[rank] : a.w[rank]);
    for (long long i0 = seg_b + t0; i0 < seg_e; i0 += stride * kUnroll) {
      float4 acc[kUnroll];
      float4 wv[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const long long i = i0 + u * stride;
        wv[u] = (!a.plain && i < seg_e) ? wl[i] : zero;
      }
      gather_sum(i0, seg_e, acc);
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const long long i = i0 + u * stride;
        if (i < seg_e) {
          wl[i] = a.plain ? acc[u] : apply_update(wv[u], acc[u], i < a.n_params4, a.scale);
          if (a.gz != nullptr) reinterpret_cast<float4*>(a.gz)[i] = zero;
        }
      }
    }
  } else if (ALGO == kPull) {
    // phase 1: reduce my shard, update my own weights (local stores only)
    const long long per = (seg_e - seg_b + world - 1) / world;
    const long long s0 = min(seg_e, seg_b + per * rank), s1 = min(seg_e, s0 + per);
    float4* wl = reinterpret_cast<float4*>(a.w[rank]);
    for (long long i0 = s0 + t0; i0 < s1; i0 += stride * kUnroll) {
      float4 acc[kUnroll];
      float4 wv[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const long long i = i0 + u * stride;
        wv[u] = i < s1 ? wl[i] : zero;
      }
      gather_sum(i0, s1, acc);
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const long long i = i0 + u * stride;
        if (i < s1) wl[i] = apply_update(wv[u], acc[u], i < a.n_params4, a.scale);
      }
    }
    if (a.gz != nullptr)
      for (long long i = seg_b + t0; i < seg_e; i += stride) reinterpret_cast<float4*>(a.gz)[i] = zero;
    // phase 2: my shard is complete in my own L2 (gpu-scope fence), tell the peers, then pull
    // every other owner's shard of the new weights.
    __threadfence();
    rank_barrier<false>(a, rank, &s_epoch);
    // One flat loop over "every element that is not mine", 4 remote loads in flight per thread;
    // consecutive strides of the grid hit different owners, so all links are busy at once.
    const long long others = per * (world - 1);
    for (long long j0 = t0; j0 < others; j0 += stride * 4) {
      float4 v[4];
      long long idx[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const long long j = j0 + u * stride;
        const int q = 1 + static_cast<int>(j / per);
        const int p = (rank + q) % world;            // stagger the owners across ranks
        idx[u] = seg_b + per * p + (j - (long long)(q - 1) * per);
        if (j < others && idx[u] < seg_e)
          v[u] = ld_nc_na_f4(reinterpret_cast<const float4*>(a.w[p]) + idx[u]);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const long long j = j0 + u * stride;
        if (j < others && idx[u] < seg_e) wl[idx[u]] = v[u];
      }
    }
  } else {
    const long long per = (seg_e - seg_b + world - 1) / world;
    const long long s0 = min(seg_e, seg_b + per * rank), s1 = min(seg_e, s0 + per);
    for (long long i0 = s0 + t0; i0 < s1; i0 += stride * kUnroll) {
      float4 acc[kUnroll];
      float4 wv[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const long long i = i0 + u * stride;
        wv[u] = (!a.plain && i < s1) ? reinterpret_cast<const float4*>(a.w[rank])[i] : zero;
      }
      if (ALGO == kNvls) {
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
          const long long i = i0 + u * stride;
          acc[u] = i < s1 ? multimem_ld_reduce_add_f4(reinterpret_cast<const float4*>(a.g_mc) + i) : zero;
        }
      } else {
        gather_sum(i0, s1, acc);
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const long long i = i0 + u * stride;
        if (i >= s1) continue;
        const float4 v = a.plain ? acc[u] : apply_update(wv[u], acc[u], i < a.n_params4, a.scale);
        if (ALGO == kNvls) {
          multimem_st_f4(reinterpret_cast<float4*>(a.plain ? a.g_mc : a.w_mc) + i, v);
        } else {
#pragma unroll
          for (int p = 0; p < kMaxRanks; ++p)
            if (p < world) st_na_f4(reinterpret_cast<float4*>(a.plain ? a.g[p] : a.w[p]) + i, v);
        }
      }
    }
  }
  }  // segments

  // With double-buffered gradients the buffer just reduced is not touched again until the step
  // after next, whose own barrier A proves every peer has finished reading it; the other buffer
  // (consumed one step ago) is what gets zeroed.  One-shot then needs no second barrier at all:
  // nobody writes remote memory.  Two-shot / NVLS still need B so that every peer's weight writes
  // have landed before the next forward pass reads W.
  if (dbg != nullptr && threadIdx.x == 0) {
    const long long t_end = ar_time_ns();
    dbg[0] += 1;
    dbg[1] += s_t[1] - s_t[0];
    dbg[2] += s_t[2] - s_t[1];
    dbg[3] += s_t[3] - s_t[2];
    dbg[4] += t_end - s_t[3];
  }
  if (a.gz != nullptr) {
    if (ALGO == kOneShot || ALGO == kPull) return;
    float4* gzl = reinterpret_cast<float4*>(a.gz);
    for (int sg = 0; sg < a.nseg; ++sg)
      for (long long i = a.seg_begin4[sg] + t0; i < a.seg_end4[sg]; i += stride) gzl[i] = zero;
    rank_barrier<true>(a, rank, &s_epoch);  // B: all writes of W have completed
    return;
  }

  rank_barrier<true>(a, rank, &s_epoch);  // B: all reads of G and all writes of W have completed

  if (!a.plain) {
    // zero the stripes of my G that block b of some rank has just consumed
    float4* gl = reinterpret_cast<float4*>(a.g[rank]);
    for (int sg = 0; sg < a.nseg; ++sg) {
      const long long seg_b = a.seg_begin4[sg], seg_e = a.seg_end4[sg];
      if (ALGO == kOneShot) {
        for (long long i = seg_b + t0; i < seg_e; i += stride) gl[i] = zero;
      } else {
        const long long per = (seg_e - seg_b + world - 1) / world;
        for (int q = 0; q < world; ++q) {
          const long long s0 = min(seg_e, seg_b + per * q), s1 = min(seg_e, s0 + per);
          for (long long i = s0 + t0; i < s1; i += stride) gl[i] = zero;
        }
      }
    }
  }
}

}  // namespace b2d

using namespace b2d;

// b2d_allreduce_desc is declared in include/b200dist.h

extern "C" size_t b2d_allreduce_signal_bytes(void) {
  return sizeof(uint32_t) * kMaxArBlocks * kMaxRanks;
}
extern "C" size_t b2d_allreduce_epoch_bytes(void) { return sizeof(uint32_t) * kMaxArBlocks; }

// Size-based dispatch: below the crossover the single pass of one-shot wins on latency; above
// it the (N-1)/N traffic of two-shot (or S/N of NVLS) wins on bandwidth.
extern "C" int b2d_allreduce_pick_algo(int64_t n_total_floats, int world, int have_multicast) {
  const int64_t bytes = n_total_floats * 4;
  // Measured on 8 x B200 (profiles/r01_allreduce.md): one-shot (single barrier, reads only) wins
  // at 2 ranks and for tiny messages; with multicast bound the switch reduction wins from
  // ~256 KB up at >= 4 ranks; without it the pull variant is the latency choice up to a few MB.
  if (world <= 2 || bytes <= (256 << 10) / world * 2) return kOneShot;
  if (have_multicast && bytes >= (256 << 10)) return kNvls;
  if (bytes <= (4 << 20)) return kPull;
  return kTwoShot;
}

extern "C" int b2d_allreduce_sgd(const b2d_allreduce_desc* d, void* stream) {
  B2D_REQUIRE(d != nullptr, "allreduce_sgd: null descriptor");
  B2D_REQUIRE(d->world >= 1 && d->world <= kMaxRanks, "allreduce_sgd: world must be in [1, %d]", kMaxRanks);
  B2D_REQUIRE(d->n_total % 4 == 0 && d->n_params % 4 == 0 && d->n_params <= d->n_total,
              "allreduce_sgd: sizes must be multiples of 4 floats");
  int algo = d->algo;
  if (algo < 0) algo = b2d_allreduce_pick_algo(d->n_total, d->world, d->g_mc != nullptr);
  B2D_REQUIRE(algo != kNvls || (d->g_mc != nullptr && (d->plain || d->w_mc != nullptr)),
              "allreduce_sgd: NVLS requested but no multicast mapping is bound");
  ArArgs a{};
  for (int p = 0; p < d->world; ++p) {
    a.w[p] = static_cast<float*>(d->w[p]);
    a.g[p] = static_cast<float*>(d->g[p]);
    a.sig[p] = static_cast<uint32_t*>(d->sig[p]);
    a.epoch[p] = static_cast<uint32_t*>(d->epoch[p]);
    a.out[p] = static_cast<float*>(d->out[p]);
  }
  a.w_mc = static_cast<float*>(d->w_mc), a.g_mc = static_cast<float*>(d->g_mc);
  a.n_params4 = d->n_params / 4, a.n_total4 = d->n_total / 4;
  a.scale = d->lr / static_cast<float>(d->world);
  a.rank = d->rank, a.world = d->world, a.emulate = d->emulate, a.plain = d->plain;
  a.gz = d->emulate ? nullptr : static_cast<float*>(d->g_zero);
  a.dbg = static_cast<long long*>(d->timeline);
  long long span4 = 0;
  if (d->nseg <= 0) {
    a.nseg = 1, a.seg_begin4[0] = 0, a.seg_end4[0] = a.n_total4;
    span4 = a.n_total4;
  } else {
    B2D_REQUIRE(d->nseg <= 2, "allreduce_sgd: at most 2 segments per launch");
    a.nseg = d->nseg;
    for (int k = 0; k < d->nseg; ++k) {
      B2D_REQUIRE(d->seg_begin[k] % 4 == 0 && d->seg_end[k] % 4 == 0 && d->seg_begin[k] <= d->seg_end[k] &&
                      d->seg_end[k] <= d->n_total,
                  "allreduce_sgd: bad segment [%lld, %lld)", (long long)d->seg_begin[k], (long long)d->seg_end[k]);
      a.seg_begin4[k] = d->seg_begin[k] / 4, a.seg_end4[k] = d->seg_end[k] / 4;
      span4 = std::max(span4, a.seg_end4[k] - a.seg_begin4[k]);
    }
  }
  if (algo < 0 || d->algo < 0) algo = b2d_allreduce_pick_algo(span4 * 4, d->world, d->g_mc != nullptr);

  // grid sized by the largest phase: one-shot and the pull phase touch the whole span per rank
  const long long work4 = (algo == kOneShot || algo == kPull) ? span4 : (span4 + d->world - 1) / d->world;
  int blocks = d->blocks > 0 ? d->blocks
                             : static_cast<int>(std::min<long long>(
                                   kMaxArBlocks, std::max<long long>(1, (work4 + kArThreads * 2 - 1) / (kArThreads * 2))));  // ~2 float4 per thread
  blocks = std::min(blocks, kMaxArBlocks);
  void (*kern)(const ArArgs) = algo == kOneShot  ? allreduce_sgd_kernel<kOneShot>
                               : algo == kTwoShot ? allreduce_sgd_kernel<kTwoShot>
                               : algo == kPull    ? allreduce_sgd_kernel<kPull>
                                                  : allreduce_sgd_kernel<kNvls>;
  if (d->emulate) blocks = std::max(1, std::min(blocks, 128 / d->world));  // cooperative: all co-resident
  B2D_REQUIRE(!(algo == kPull && d->plain), "allreduce_sgd: the pull algorithm is fused-SGD only");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (d->emulate) {
    // All ranks in one launch: the blocks wait on one another, so they must be co-resident.
    void* params[] = {&a};
    B2D_CUDA_CHECK(cudaLaunchCooperativeKernel(reinterpret_cast<const void*>(kern), dim3(blocks, d->world),
                                               dim3(kArThreads), params, 0, st));
    count_launch();
  } else {
    B2D_CUDA_CHECK(launch_pdl(kern, dim3(blocks), dim3(kArThreads), 0, st, a));
  }
  return B2D_OK;
}
