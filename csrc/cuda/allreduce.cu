// allreduce.cu — the communication-bound hot path: cross-replica gradient all-reduce FUSED with
// the 1/num_workers averaging and the SGD update, over NVLink peer memory (sm_100a).
//
// Every rank holds a weight arena W_r and a gradient arena G_r in symmetric memory: every GPU
// can load/store every peer's arenas directly through NVSwitch (csrc/runtime/symm.cpp).
// These kernels replace "batch_all_reduce: 6 all-reduces" + 4 metric all-reduces + the optimizer
// (README.md:403-407, 302).  Variants of allreduce_sgd_kernel, chosen per message size
// (b2d_allreduce_pick_algo) or by the producer of the gradients:
//
//   one-shot   every rank reads all peers' G, sums in rank order (bit-identical everywhere),
//              updates its own W.                          traffic/GPU: (N-1) S in
//   two-shot   rank r owns shard r: reads shard r of every G, averages, applies SGD, and writes
//              the *updated weights* into every peer's W — the all-gather carries weights, so
//              no optimizer pass exists at all.            traffic/GPU: (N-1)/N S in + out
//   pull       two-shot without remote writes: rank r reduces shard r into its OWN weights, then
//              (release / acquire flag pair) every rank pulls the other shards — block b pulls
//              exactly what the owner's block b wrote.  Only used when asked for by name.
//   NVLS       like two-shot, but the switch does the sum (multimem.ld_reduce) and the
//              broadcast (multimem.st).                     traffic/GPU: S/N in + S/N out
//   scattered  the producing GEMMs already added their tiles into the owners' arenas
//              (gemm_tcgen05.cu, GEMM -> reduce-scatter): the owner reads its own shard, applies
//              SGD, broadcasts the weights.                 traffic/GPU: S/N out (+ the GEMM's reds)
// plus ll_allreduce_sgd_kernel, the push exchange for small buckets (below).
//
// Elements >= n_params are the metric tail slots (loss sum, correct count, ...): they are
// summed, not averaged, and *added* to the running sums kept in W's tail.
// After the update each rank zeroes its own G, so the next backward can accumulate into it.
//
// Cross-GPU barriers are per block: block b of every rank meets block b of every other rank
// on monotone epoch counters in the peers' signal pads (relaxed system-scope flags where the
// barrier only announces earlier kernels' results, st.release.sys + one ld.acquire.sys where it
// publishes this kernel's remote writes).  Waits are bounded by %globaltimer.
//
// Step schedule with weight ping-pong (engine/trainer.py): step t reads W[t&1] and the update
// writes W[(t+1)&1], so an update can never collide with a kernel of the same step that still
// reads the weights.  The gradient buckets of the later layers are therefore reduced on a side
// branch of the step graph as soon as their weight gradients are final, under the rest of the
// backward pass; only the first layer's (small) bucket and the metric sums remain on the
// critical path, and they travel by the LL kernel below: every rank pushes {value, epoch} pairs
// as single 8-byte stores into the peers' inboxes, so the data is its own arrival flag (one
// one-way NVLink trip, no separate barrier).  That exchange is also the step's closing barrier:
// a peer's LL flag proves that all of its earlier kernels — including the ones that wrote into
// this rank's next weight buffer — have completed.
#include <cstdlib>

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
  float* wo[kMaxRanks];       // where the updated weights go: == w (in place) or the other weight buffer
  float* wo_mc;
  int no_exit;                // 1: skip barrier B (a later kernel of the step closes it, see above)
  int strict;                 // 1: entry barrier with st.release.sys / ld.acquire.sys
  int debug;                  // B200DIST_AR_DEBUG: random delays, flag monotonicity checks, poison
  long long timeout_ns;       // a barrier that waits longer than this traps (fail-stop)
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
__device__ __forceinline__ uint32_t ar_hash(uint32_t x) {
  x ^= x >> 16, x *= 0x7feb352dU, x ^= x >> 15, x *= 0x846ca68bU, x ^= x >> 16;
  return x;
}
// Debug mode: a pseudo-random pause per (rank, block, epoch, phase) widens every race window.
__device__ __forceinline__ void ar_debug_delay(const ArArgs& a, int rank, uint32_t epoch, int phase) {
  if (a.debug && threadIdx.x == 0) {
    const uint32_t h = ar_hash(rank * 0x9e3779b9U ^ blockIdx.x * 0x85ebca6bU ^ epoch * 0xc2b2ae35U ^ phase);
    if ((h & 3) == 0) __nanosleep(h % 20000);
  }
}

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
  ar_debug_delay(a, rank, e, 0);
  if (a.debug) __syncthreads();
  if (threadIdx.x < a.world) {
    const int p = threadIdx.x;
    if (kPublish)
      st_release_sys(a.sig[p] + b * kMaxRanks + rank, e);
    else
      st_relaxed_sys(a.sig[p] + b * kMaxRanks + rank, e);
    if (dbg != nullptr && threadIdx.x == 0) dbg[0] = ar_time_ns();   // signals sent (dbg = &local[2])
    const uint32_t* mine = a.sig[rank] + b * kMaxRanks + p;
    long long t_start = 0;
    uint32_t spins = 0, v;
    // Poll relaxed (an acquire load per poll made this wait 4-8 us longer on B200); where the barrier
    // publishes remote writes, ONE acquire load of the satisfied flag orders the later accesses behind it.
    while (static_cast<int32_t>((v = ld_relaxed_sys(mine)) - e) < 0) {
      if ((++spins & 0x3ff) == 0) {  // a peer died: fail loudly instead of hanging the GPU forever
        const long long now = ar_time_ns();
        if (t_start == 0) t_start = now;
        if (now - t_start > a.timeout_ns) {
          printf("b200dist: rank %d block %d timed out waiting for rank %d (epoch %u, saw %u)\n", rank, b, p, e, v);
          __trap();
        }
      }
    }
    if (kPublish) v = ld_acquire_sys(mine);
    if (a.debug) {
      // Monotone epochs: a flag never runs backwards, and a peer can be at most one barrier ahead
      // of this block (it cannot pass barrier e+1 before this block has signalled it).
      uint32_t* seen = a.epoch[rank] + kMaxArBlocks + b * kMaxRanks + p;   // behind the epochs, same allocation
      if (static_cast<int32_t>(v - *seen) < 0 || static_cast<int32_t>(v - e) > 1) {
        printf("b200dist[debug]: rank %d block %d: flag of rank %d went %u -> %u at epoch %u\n", rank, b, p, *seen, v, e);
        __trap();
      }
      *seen = v;
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

enum ArAlgo : int { kOneShot = 0, kTwoShot = 1, kNvls = 2, kPull = 3, kScattered = 4 };

// kW = compile-time bound on the number of ranks (2, 4 or 8): the gather keeps kW x kUnroll float4 in
// registers, and the kernel has to stay light — its blocks share the SMs with the backward kernels they
// overlap (two 512-thread blocks per SM must fit next to a GEMM CTA, hence at most 64 registers when
// kW <= 4; a block that has to wait for a free SM shows up as a late barrier arrival on every peer).
template <int ALGO, int kW>
__global__ void __launch_bounds__(kArThreads, kW <= 4 ? 2 : 1) allreduce_sgd_kernel(const ArArgs a) {
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
  // A: every replica's gradients are complete
  if (a.strict)
    rank_barrier<true>(a, rank, &s_epoch, dbg == nullptr ? nullptr : &s_t[2]);
  else
    rank_barrier<false>(a, rank, &s_epoch, dbg == nullptr ? nullptr : &s_t[2]);
  ar_debug_delay(a, rank, s_epoch, 1);
  if (a.debug) __syncthreads();
  // Zeroing of the gradient buffer consumed one step ago.  Debug mode first poisons it and keeps the
  // poison up for a while: a peer that still read it (a protocol violation) would turn its weights into NaNs.
  auto zero_range = [&](float4* base, long long b4, long long e4) {
    if (a.debug) {
      const float4 poison = make_float4(__int_as_float(0x7fc0dead), __int_as_float(0x7fc0dead),
                                        __int_as_float(0x7fc0dead), __int_as_float(0x7fc0dead));
      for (long long i = b4 + t0; i < e4; i += stride) base[i] = poison;
      __threadfence_system();
      __nanosleep(20000);
    }
    for (long long i = b4 + t0; i < e4; i += stride) base[i] = zero;
  };
  if (dbg != nullptr && threadIdx.x == 0) s_t[3] = ar_time_ns();

  // Peer loads have ~2 us latency.  All loads of a batch — every peer x kUnroll float4 — are issued
  // before the first use, and only then summed in fixed rank order (identical bits everywhere).
  constexpr int kUnroll = 2;
  auto gather_sum = [&](long long i0, long long limit, float4 (&acc)[kUnroll]) {
    float4 v[kW][kUnroll];
#pragma unroll
    for (int p = 0; p < kW; ++p)
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
    for (int p = 0; p < kW; ++p)
      if (p < world) {
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) acc[u] = f4_add(acc[u], v[p][u]);
      }
  };
#pragma unroll 1
  for (int sg = 0; sg < a.nseg; ++sg) {
  const long long seg_b = a.seg_begin4[sg], seg_e = a.seg_end4[sg];
  if (ALGO == kOneShot) {
    const float4* wl = reinterpret_cast<const float4*>(a.w[rank]);
    float4* wol = reinterpret_cast<float4*>(a.plain ? a.out[rank] : a.wo[rank]);
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
          wol[i] = a.plain ? acc[u] : apply_update(wv[u], acc[u], i < a.n_params4, a.scale);
          if (a.gz != nullptr && !a.debug) reinterpret_cast<float4*>(a.gz)[i] = zero;
        }
      }
    }
    if (a.gz != nullptr && a.debug) zero_range(reinterpret_cast<float4*>(a.gz), seg_b, seg_e);
  } else if (ALGO == kPull) {
    // phase 1: reduce my shard, write my own copy of the new weights (local stores only)
    const long long per = (seg_e - seg_b + world - 1) / world;
    const long long s0 = min(seg_e, seg_b + per * rank), s1 = min(seg_e, s0 + per);
    const float4* wl = reinterpret_cast<const float4*>(a.w[rank]);
    float4* wol = reinterpret_cast<float4*>(a.wo[rank]);
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
        if (i < s1) wol[i] = apply_update(wv[u], acc[u], i < a.n_params4, a.scale);
      }
    }
    if (a.gz != nullptr) zero_range(reinterpret_cast<float4*>(a.gz), seg_b, seg_e);
    // phase 2: tell the peers that my shard is written, then pull every other owner's shard.
    // The barrier pairs block b of this rank with block b of the owner ONLY, so block b pulls
    // exactly the elements the owner's block b wrote in phase 1: offsets t0 + m * stride inside
    // the owner's shard (any other mapping reads elements of blocks it never synchronised with).
    rank_barrier<true>(a, rank, &s_epoch);
    ar_debug_delay(a, rank, s_epoch, 2);
    if (a.debug) __syncthreads();
    for (int q = 1; q < world; ++q) {
      const int p = (rank + q) % world;               // stagger the owners across ranks
      const long long p0 = min(seg_e, seg_b + per * p), p1 = min(seg_e, p0 + per);
      const float4* src = reinterpret_cast<const float4*>(a.wo[p]);
      for (long long i0 = p0 + t0; i0 < p1; i0 += stride * 4) {
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const long long i = i0 + u * stride;
          if (i < p1) v[u] = ld_nc_na_f4(src + i);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const long long i = i0 + u * stride;
          if (i < p1) wol[i] = v[u];
        }
      }
    }
  } else if (ALGO == kScattered) {
    // The producing GEMMs of all replicas have already added their results into the owners' arenas
    // (gemm_tcgen05.cu, kFlagScatter): my shard of my own G is the complete sum.  Read it locally, zero it
    // in place for the step after next, update, and broadcast the new weights.
    const long long per = (seg_e - seg_b + world - 1) / world;
    const long long s0 = min(seg_e, seg_b + per * rank), s1 = min(seg_e, s0 + per);
    float4* gl = reinterpret_cast<float4*>(a.g[rank]);
    const float4* wl = reinterpret_cast<const float4*>(a.w[rank]);
    for (long long i0 = s0 + t0; i0 < s1; i0 += stride * 4) {
      float4 acc[4], wv[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const long long i = i0 + u * stride;
        if (i < s1) acc[u] = __ldcg(gl + i), wv[u] = wl[i];
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const long long i = i0 + u * stride;
        if (i >= s1) continue;
        gl[i] = zero;
        const float4 v = apply_update(wv[u], acc[u], i < a.n_params4, a.scale);
        if (a.wo_mc != nullptr) {
          multimem_st_f4(reinterpret_cast<float4*>(a.wo_mc) + i, v);
        } else {
#pragma unroll
          for (int p = 0; p < kW; ++p)
            if (p < world) st_na_f4(reinterpret_cast<float4*>(a.wo[p]) + i, v);
        }
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
          multimem_st_f4(reinterpret_cast<float4*>(a.plain ? a.g_mc : a.wo_mc) + i, v);
        } else {
#pragma unroll
          for (int p = 0; p < kW; ++p)
            if (p < world) st_na_f4(reinterpret_cast<float4*>(a.plain ? a.g[p] : a.wo[p]) + i, v);
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
    dbg[5] = s_t[0], dbg[6] = s_t[3], dbg[7] = t_end;   // absolute stamps of the last launch: start, past barrier A, end
  }
  if (ALGO == kScattered) {
    // in-place zeroing is done; the other gradient buffer's range was zeroed the same way one step ago
    if (!a.no_exit) rank_barrier<true>(a, rank, &s_epoch);
    return;
  }
  if (a.gz != nullptr) {
    if (ALGO == kOneShot || ALGO == kPull) return;
    for (int sg = 0; sg < a.nseg; ++sg)
      zero_range(reinterpret_cast<float4*>(a.gz), a.seg_begin4[sg], a.seg_end4[sg]);
    // B: all writes of W have completed.  Skipped when a later kernel of the same step closes the
    // step for every bucket at once (the LL exchange of the last bucket).
    if (!a.no_exit) rank_barrier<true>(a, rank, &s_epoch);
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

// ----------------------------------------------------------------------------------------
// LL all-reduce + SGD for small buckets: the data is its own flag.
//
// Every rank pushes each of its gradient values to every peer as ONE 8-byte store
// {value bits, epoch}; 8-byte aligned stores are single-copy atomic, so a reader that sees the
// epoch also sees the value — no fence, no separate barrier, one one-way NVLink trip.  Each rank
// then sums the world contributions per element in rank order (bit-identical everywhere) and
// applies the update to its own weights only (no remote weight writes).  Inboxes are double
// buffered by epoch parity, so a fast peer's next exchange cannot overwrite values that a slow
// rank has not read yet.  Up to kLlMaxFloats values per launch (first layer + metric tail).
// ----------------------------------------------------------------------------------------
constexpr int kLlThreads = 256;
constexpr int kLlMaxFloats = 16384;
constexpr int kLlMaxBlocks = kLlMaxFloats / kLlThreads;

struct LlArgs {
  unsigned long long* inbox[kMaxRanks];  // [2 parities][kMaxRanks senders][kLlMaxFloats] per rank
  uint32_t* epoch[kMaxRanks];            // per-block launch counters (local to each rank)
  const float* w[kMaxRanks];             // weights read (only [rank]; all set in emulation)
  float* wo[kMaxRanks];                  // weights written
  const float* g[kMaxRanks];             // gradients of this step
  float* gz[kMaxRanks];                  // the other gradient buffer: zeroed (may be null)
  long long seg_begin[2], seg_end[2];    // float ranges of the arenas
  int nseg;
  long long n_params;
  float scale;
  int rank, world, emulate, debug;
  long long timeout_ns;
  long long* dbg;                        // optional: 4 absolute globaltimer stamps per block (last launch)
};

__device__ __forceinline__ void st_relaxed_sys_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_sys_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

__global__ void __launch_bounds__(kLlThreads) ll_allreduce_sgd_kernel(const LlArgs a) {
  __shared__ uint32_t s_epoch;
  const int rank = a.emulate ? blockIdx.y : a.rank;
  const int world = a.world;
  long long* dbg = a.dbg == nullptr ? nullptr : a.dbg + 4 * blockIdx.x;
  if (dbg != nullptr && threadIdx.x == 0) dbg[0] = ar_time_ns();
  // Everything up to the gradient load is independent of the kernel this one is launched behind (the epoch
  // counter is private to this kernel sequence, the weights were written by the previous step's update): with
  // programmatic dependent launch it runs under the tail of the last backward kernel.
  if (threadIdx.x == 0) {
    uint32_t* ep = a.epoch[rank] + blockIdx.x;
    uint32_t e = *ep + 1;
    if (e == 0) e = 1;               // 0 is the "never written" value of the inboxes
    *ep = e;
    s_epoch = e;
  }
  __syncthreads();
  const uint32_t e = s_epoch;
  const long long len0 = a.seg_end[0] - a.seg_begin[0];
  const long long n = len0 + (a.nseg > 1 ? a.seg_end[1] - a.seg_begin[1] : 0);
  const long long i = (long long)blockIdx.x * kLlThreads + threadIdx.x;
  if (a.debug && threadIdx.x == 0) {
    const uint32_t h = ar_hash(rank * 0x9e3779b9U ^ blockIdx.x * 0x85ebca6bU ^ e * 0xc2b2ae35U ^ 7);
    if ((h & 3) == 0) __nanosleep(h % 20000);
  }
  if (a.debug) __syncthreads();
  const bool active = i < n;
  const long long idx = !active ? 0 : (i < len0 ? a.seg_begin[0] + i : a.seg_begin[1] + (i - len0));
  const size_t slot = static_cast<size_t>(e & 1) * kMaxRanks * kLlMaxFloats + i;
  const float wv = active ? a.w[rank][idx] : 0.f;
  pdl_wait();     // this replica's backward has finished writing these gradients
  pdl_trigger();
  if (dbg != nullptr && threadIdx.x == 0) dbg[1] = ar_time_ns();
  if (!active) return;
  const float mine = a.g[rank][idx];
  const unsigned long long packed = (static_cast<unsigned long long>(e) << 32) | __float_as_uint(mine);
  for (int q = 1; q < world; ++q) {
    const int p = (rank + q) % world;
    st_relaxed_sys_u64(a.inbox[p] + slot + static_cast<size_t>(rank) * kLlMaxFloats, packed);
  }
  float sum = 0.f;
  long long t_start = 0;
  for (int p = 0; p < world; ++p) {          // fixed order: identical bits on every rank
    float v = mine;
    if (p != rank) {
      const unsigned long long* src = a.inbox[rank] + slot + static_cast<size_t>(p) * kLlMaxFloats;
      unsigned long long u;
      uint32_t spins = 0;
      while (static_cast<uint32_t>((u = ld_relaxed_sys_u64(src)) >> 32) != e) {
        if ((++spins & 0x3ff) == 0) {
          const long long now = ar_time_ns();
          if (t_start == 0) t_start = now;
          if (now - t_start > a.timeout_ns) {
            printf("b200dist: rank %d timed out waiting for the LL value %lld of rank %d (epoch %u, saw %u)\n",
                   rank, i, p, e, static_cast<uint32_t>(u >> 32));
            __trap();
          }
        }
      }
      v = __uint_as_float(static_cast<uint32_t>(u));
    }
    sum += v;
  }
  a.wo[rank][idx] = idx < a.n_params ? wv - a.scale * sum : wv + sum;
  if (a.gz[rank] != nullptr) a.gz[rank][idx] = 0.f;
  if (dbg != nullptr && threadIdx.x == 0) dbg[2] = ar_time_ns();
}

}  // namespace b2d

using namespace b2d;

// b2d_allreduce_desc is declared in include/b200dist.h

static bool ar_env_flag(const char* name) {
  const char* v = getenv(name);
  return v != nullptr && v[0] != '\0' && v[0] != '0';
}
// B200DIST_AR_TIMEOUT_S: how long a cross-GPU wait may last before the kernel traps (default 30 s).
static long long ar_timeout_ns() {
  const char* v = getenv("B200DIST_AR_TIMEOUT_S");
  double s = v != nullptr ? atof(v) : 30.0;
  if (!(s > 0.0)) s = 30.0;
  return static_cast<long long>(s * 1e9);
}
extern "C" size_t b2d_allreduce_signal_bytes(void) {
  return sizeof(uint32_t) * kMaxArBlocks * kMaxRanks;
}
// per-block epochs, followed by the debug mode's "last flag seen" words per (block, peer)
extern "C" size_t b2d_allreduce_epoch_bytes(void) { return sizeof(uint32_t) * kMaxArBlocks * (1 + kMaxRanks); }

// Size-based dispatch: below the crossover the single pass of one-shot wins on latency; above
// it the (N-1)/N traffic of two-shot (or S/N of NVLS) wins on bandwidth.
extern "C" int b2d_allreduce_pick_algo(int64_t n_total_floats, int world, int have_multicast) {
  const int64_t bytes = n_total_floats * 4;
  // Measured on 2 / 4 / 8 x B200 (profiles/r02_allreduce.md):
  // * 2 ranks: one-shot and two-shot move the same bytes, two-shot uses both link directions and wins from
  //   a few MB up (16 MB: 45 vs 46 us, 1 GB: 2.09 ms); NVLS never wins with two ranks.
  // * with multicast bound and >= 3 ranks the switch-side reduction is the fastest variant from 8 KB up
  //   (8 ranks: 64 KB 15.7 vs 20.6 us two-shot, 16 MB 70 vs 82 us, 256 MB 708 vs 897 us); below that one
  //   barrier less makes one-shot as good.
  // * without multicast two-shot wins from 16 KB; the pull variant (an acquire/release pair per phase) is
  //   never the fastest and is only used when asked for by name.
  if (world <= 2) return bytes >= (4 << 20) ? kTwoShot : kOneShot;
  if (have_multicast) return bytes >= (8 << 10) ? kNvls : kOneShot;
  return bytes >= (16 << 10) ? kTwoShot : kOneShot;
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
  for (int p = 0; p < d->world; ++p) a.wo[p] = d->w_out[p] != nullptr ? static_cast<float*>(d->w_out[p]) : a.w[p];
  a.wo_mc = d->w_out_mc != nullptr ? static_cast<float*>(d->w_out_mc) : a.w_mc;
  a.no_exit = (d->flags & B2D_AR_NO_EXIT_BARRIER) != 0 && a.gz != nullptr;
  a.strict = (d->flags & B2D_AR_STRICT_ENTRY) != 0 || ar_env_flag("B200DIST_AR_STRICT");
  a.debug = ar_env_flag("B200DIST_AR_DEBUG");
  a.timeout_ns = ar_timeout_ns();
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
  void (*kern)(const ArArgs) = nullptr;
#define B2D_AR_PICK(W)                                                                    \
  kern = algo == kOneShot    ? allreduce_sgd_kernel<kOneShot, W>                          \
         : algo == kTwoShot   ? allreduce_sgd_kernel<kTwoShot, W>                         \
         : algo == kPull      ? allreduce_sgd_kernel<kPull, W>                            \
         : algo == kScattered ? allreduce_sgd_kernel<kScattered, W>                       \
                              : allreduce_sgd_kernel<kNvls, W>
  if (d->world <= 2) {
    B2D_AR_PICK(2);
  } else if (d->world <= 4) {
    B2D_AR_PICK(4);
  } else {
    B2D_AR_PICK(8);
  }
#undef B2D_AR_PICK
  if (d->emulate) blocks = std::max(1, std::min(blocks, 128 / d->world));  // cooperative: all co-resident
  B2D_REQUIRE(!((algo == kPull || algo == kScattered) && d->plain),
              "allreduce_sgd: the pull and scattered algorithms are fused-SGD only");
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

extern "C" size_t b2d_ll_inbox_bytes(void) {
  return sizeof(unsigned long long) * 2 * kMaxRanks * kLlMaxFloats;
}
extern "C" size_t b2d_ll_epoch_bytes(void) { return sizeof(uint32_t) * kLlMaxBlocks; }
extern "C" int64_t b2d_ll_max_floats(void) { return kLlMaxFloats; }

extern "C" int b2d_ll_allreduce_sgd(const b2d_ll_desc* d, void* stream) {
  B2D_REQUIRE(d != nullptr, "ll_allreduce_sgd: null descriptor");
  B2D_REQUIRE(d->world >= 1 && d->world <= kMaxRanks, "ll_allreduce_sgd: world must be in [1, %d]", kMaxRanks);
  B2D_REQUIRE(d->nseg >= 1 && d->nseg <= 2, "ll_allreduce_sgd: one or two segments");
  LlArgs a{};
  long long n = 0;
  for (int k = 0; k < d->nseg; ++k) {
    B2D_REQUIRE(d->seg_begin[k] >= 0 && d->seg_begin[k] <= d->seg_end[k],
                "ll_allreduce_sgd: bad segment [%lld, %lld)", (long long)d->seg_begin[k], (long long)d->seg_end[k]);
    a.seg_begin[k] = d->seg_begin[k], a.seg_end[k] = d->seg_end[k];
    n += d->seg_end[k] - d->seg_begin[k];
  }
  B2D_REQUIRE(n >= 1 && n <= kLlMaxFloats, "ll_allreduce_sgd: %lld values per launch, the limit is %d", n, kLlMaxFloats);
  a.nseg = d->nseg;
  for (int p = 0; p < d->world; ++p) {
    a.inbox[p] = static_cast<unsigned long long*>(d->inbox[p]);
    a.epoch[p] = static_cast<uint32_t*>(d->epoch[p]);
    a.w[p] = static_cast<const float*>(d->w[p]);
    a.wo[p] = d->w_out[p] != nullptr ? static_cast<float*>(d->w_out[p]) : static_cast<float*>(d->w[p]);
    a.g[p] = static_cast<const float*>(d->g[p]);
    a.gz[p] = static_cast<float*>(d->g_zero[p]);
  }
  a.n_params = d->n_params;
  a.scale = d->lr / static_cast<float>(d->world);
  a.rank = d->rank, a.world = d->world, a.emulate = d->emulate;
  a.debug = ar_env_flag("B200DIST_AR_DEBUG");
  a.timeout_ns = ar_timeout_ns();
  a.dbg = static_cast<long long*>(d->timeline);
  const int blocks = static_cast<int>((n + kLlThreads - 1) / kLlThreads);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (d->emulate) {
    void* params[] = {&a};
    B2D_CUDA_CHECK(cudaLaunchCooperativeKernel(reinterpret_cast<const void*>(ll_allreduce_sgd_kernel),
                                               dim3(blocks, d->world), dim3(kLlThreads), params, 0, st));
    count_launch();
  } else {
    B2D_CUDA_CHECK(launch_pdl(ll_allreduce_sgd_kernel, dim3(blocks), dim3(kLlThreads), 0, st, a));
  }
  return B2D_OK;
}
