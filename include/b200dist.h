/* b200dist.h — C ABI of the native libraries of distributed_b200.
 *
 * Two shared objects are built in-tree (distributed_b200/_native/build.py):
 *   libb200dist_cuda.so  sm_100a kernels + symmetric-memory runtime   (functions b2d_*)
 *   libb200dist_cpu.so   C++ reference ops + shared-memory all-reduce  (functions b2c_*)
 *
 * All entry points return 0 on success and a negative b2d_status (or a positive cudaError_t /
 * CUresult, as noted) on failure; b2d_last_error() returns a human-readable message for the
 * calling thread.  Streams are passed as void* (cudaStream_t).  Unless stated otherwise the
 * functions are asynchronous with respect to the host and allocate nothing.
 *
 * The reference (Mrhs121/distributed) contains no native code; the citations below point at
 * the README line whose behaviour the kernel implements.
 */
#ifndef B200DIST_H_
#define B200DIST_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum b2d_status {
  B2D_OK = 0,
  B2D_ERR_INVALID = -1,     /* bad argument (shape, alignment, dtype)            */
  B2D_ERR_CUDA = -2,        /* a CUDA runtime / driver call failed               */
  B2D_ERR_UNSUPPORTED = -3, /* feature not available on this device / build      */
  B2D_ERR_LIMIT = -4,       /* size beyond a documented compile-time maximum     */
  B2D_ERR_SYS = -5          /* an OS call (socket, mmap, ...) failed             */
};

const char* b2d_last_error(void);
/* Number of kernels launched by this library on the calling process since load. */
uint64_t b2d_launch_count(void);

enum b2d_dtype { B2D_F32 = 0, B2D_BF16 = 1 };

/* ---------------------------------------------------------------------------------------
 * tcgen05 GEMM core (Dense fwd/dgrad/wgrad; README.md:296-297).
 *   D[M,N] = epilogue( A[M,K] * B[K,N] )
 * Operands are described by their memory layout:
 *   a_mn_major = 0 : A stored [M][lda]  (K contiguous)      1 : A stored [K][lda] (M contiguous)
 *   b_mn_major = 0 : B stored [N][ldb]  (K contiguous)      1 : B stored [K][ldb] (N contiguous)
 * dtype F32 runs kind::tf32 MMAs on fp32 storage; BF16 runs kind::f16.  Accumulation is fp32
 * in tensor memory.  Leading dimensions are in elements and must make rows 16-byte aligned.
 * Epilogue: (+bias[N]) -> (relu) -> (zero where mask[m][n] <= 0) -> store | atomic accumulate.
 * splits > 1 partitions K across CTAs; with accumulate == 0 the partial sums meet in a
 * caller-provided fp32 workspace and the last CTA of a tile applies the epilogue (the
 * workspace and counters must be zero before the first use and are left zero afterwards).
 * ------------------------------------------------------------------------------------- */
typedef struct b2d_gemm_desc {
  int dtype;     /* b2d_dtype of A and B                                   */
  int out_dtype; /* b2d_dtype of D                                         */
  int M, N, K;
  const void* A;
  int64_t lda;
  int a_mn_major;
  const void* B;
  int64_t ldb;
  int b_mn_major;
  void* D;
  int64_t ldd;
  const float* bias;  /* optional, [N]                                     */
  const void* mask;   /* optional, [M][ldmask] of mask_dtype: D = 0 where mask <= 0 */
  int64_t ldmask;
  int relu;
  int accumulate;     /* 1: D += result (fp32 atomics; D must be fp32)     */
  int splits;         /* 0 = choose automatically                          */
  float* workspace;   /* split-K scratch, b2d_gemm_workspace_bytes()       */
  uint32_t* counters; /* split-K tile counters, b2d_gemm_counter_count()   */
  int mask_dtype;     /* b2d_dtype of mask (0 = fp32)                      */
} b2d_gemm_desc;

typedef struct b2d_gemm_plan b2d_gemm_plan;

/* Resolved launch geometry for a descriptor (splits after clamping etc.). */
typedef struct b2d_gemm_geometry {
  int block_n, m_tiles, n_tiles, splits, k_blocks, stages, smem_bytes;
  int ctas; /* persistent CTAs launched: min(m_tiles * n_tiles * splits, SMs x CTAs that fit an SM together) */
  int pair; /* 1: CTA-pair kernel (tcgen05 cta_group::2): 256 x block_n tiles, m_tiles counts 256-row tiles  */
} b2d_gemm_geometry;

int b2d_gemm_query(const b2d_gemm_desc* desc, b2d_gemm_geometry* out);
size_t b2d_gemm_workspace_bytes(const b2d_gemm_desc* desc);
size_t b2d_gemm_counter_count(const b2d_gemm_desc* desc);
/* Encodes the TMA descriptors (host only, no device allocation). */
int b2d_gemm_plan_create(const b2d_gemm_desc* desc, b2d_gemm_plan** out);
int b2d_gemm_plan_run(const b2d_gemm_plan* plan, void* stream);
void b2d_gemm_plan_destroy(b2d_gemm_plan* plan);
/* GEMM -> reduce-scatter in one kernel (the weight-gradient GEMM followed by the gradient all-reduce of
 * README.md:403): an accumulating fp32 plan whose D is a range of the replicated gradient arena adds every
 * 16-byte vector of its result into the arena of the replica that owns that vector's shard — directly over
 * NVLink from the epilogue — instead of into the local arena.  The shard of an element at arena offset i is
 * min(world - 1, (i - seg_begin) / per); b2d_allreduce_sgd with algo 4 then finishes the update from the
 * owner's complete sums. */
typedef struct b2d_gemm_scatter {
  void* base[8];      /* gradient arena of every replica as mapped in this process   */
  int64_t d_off;      /* float offset of D inside the arena                          */
  int64_t seg_begin;  /* float offset of the sharded range                           */
  int64_t per;        /* floats per shard (multiple of 4)                            */
  int rank, world;
} b2d_gemm_scatter;
int b2d_gemm_plan_set_scatter(b2d_gemm_plan* plan, const b2d_gemm_scatter* sc);
/* Latency analysis (tools/gemm_timeline.py): attach a buffer of 16 x int64 per persistent CTA (geometry.ctas)
 * that receives %globaltimer stamps — start, prologue done, upstream kernel complete, first tiles landed, first
 * tile's MMAs issued, accumulator visible to the epilogue, epilogue done, end — or detach with NULL. */
void b2d_gemm_plan_set_timeline(b2d_gemm_plan* plan, long long* buf);

/* The same GEMM for callers that cannot allocate device memory (a host that has given the
 * whole GPU to its own pool, a plug-in inside another runtime's callback, a latency-critical
 * loop that cannot take an allocation's device-wide synchronisation).  desc->workspace and
 * desc->counters are ignored: the split-K scratch is a fixed array inside the library
 * (b2d_gemm_noalloc_max_workspace_floats() partial sums = 1 MiB, b2d_gemm_noalloc_max_tiles()
 * output tiles), present on every GPU that loads the library until the process exits.
 * Problems whose rendezvous needs more return B2D_ERR_LIMIT.  The call allocates nothing, on
 * host or device.  All calls on a device share the array: each clears what it will use before
 * its kernel, but calls that overlap on the GPU race.  The first call on a device resolves the
 * array's address; make it with that device current and before any stream capture.
 * This entry point has no counterpart in the reference. */
int b2d_gemm_noalloc(const b2d_gemm_desc* desc, void* stream);
size_t b2d_gemm_noalloc_max_workspace_floats(void);
size_t b2d_gemm_noalloc_max_tiles(void);

/* ---------------------------------------------------------------------------------------
 * Implicit-GEMM convolution on the same tensor-core kernel (Conv2D, README.md:293).  Forward and
 * filter gradient take any stride (TMA traversal stride); the data gradient is unit-stride.
 * forward: x [N,H,W,Ci], w [KH,KW,Ci,Co] -> y [N,OH,OW,Co] (+bias, ReLU)
 * dgrad  : x is the OUTPUT dx, y is the INPUT dy; optional ReLU mask shaped like dx
 * wgrad  : w is the OUTPUT dW (fp32, accumulated), y is dy
 * b2d_conv_tc_supported() tells whether a shape qualifies (whole 128-byte channel chunks,
 * output width <= 128); otherwise use the direct kernels below.  fwd/dgrad plans are
 * b2d_gemm_plan objects (run/destroy with the GEMM functions).
 * ------------------------------------------------------------------------------------- */
typedef struct b2d_conv_desc {
  int dtype, out_dtype;
  int N, H, W, Ci, Co, KH, KW, pad_h, pad_w;
  const void* x;
  const void* w;
  void* y;
  const float* bias;
  const void* mask; /* optional ReLU mask of the output (dgrad): same shape as the output, mask_dtype */
  int relu;
  int stride;   /* forward / wgrad: input step per output pixel (0 or 1 = unit stride)          */
  int OH, OW;   /* output size; 0 = derive from symmetric padding (TF 'same' may be asymmetric:  */
                /* pad_h / pad_w are the top / left amounts, the output size fixes the rest)    */
  int mask_dtype;
  const void* addend; /* unit-stride dgrad only: dx = mask(conv_transpose(dy, w) + addend), addend like dx */
} b2d_conv_desc;
typedef struct b2d_conv_wgrad_plan b2d_conv_wgrad_plan;
int b2d_conv_tc_supported(const b2d_conv_desc* desc, int dgrad);
int b2d_conv_fwd_plan_create(const b2d_conv_desc* desc, b2d_gemm_plan** out);
int b2d_conv_dgrad_plan_create(const b2d_conv_desc* desc, b2d_gemm_plan** out);
/* Strided data gradient: one plan per output-pixel parity class (a, b) in [0, stride)^2; *out is
 * NULL for a class without filter taps (clear dx first with b2d_memset_async). */
int b2d_conv_dgrad_class_plan_create(const b2d_conv_desc* desc, int a, int b, b2d_gemm_plan** out);
int b2d_memset_async(void* ptr, int value, size_t bytes, void* stream);
int b2d_conv_wgrad_plan_create(const b2d_conv_desc* desc, b2d_conv_wgrad_plan** out);
int b2d_conv_wgrad_plan_run(const b2d_conv_wgrad_plan* plan, void* stream);
void b2d_conv_wgrad_plan_destroy(b2d_conv_wgrad_plan* plan);

/* ---------------------------------------------------------------------------------------
 * Non-GEMM kernels (csrc/cuda/nn_kernels.cu).  fp32, NHWC / HWIO / (in,out).  One launch each.
 * ------------------------------------------------------------------------------------- */
/* Conv2D(C_in=1, KSxKS valid) + bias + ReLU + MaxPool2x2 in one kernel (README.md:293-294);
 * code = argmax position 0..3, or 4 where ReLU clamped.  KS in {3,5}. */
int b2d_conv_c1_pool_fwd(const float* x, const float* w, const float* bias, float* y, uint8_t* code,
                         int N, int H, int W, int Co, int KS, void* stream);
/* Pool-backward + ReLU-backward + filter/bias gradient of the layer above (accumulates). */
int b2d_conv_c1_pool_bwd(const float* x, const float* dy, const uint8_t* code, float* dw, float* db,
                         int N, int H, int W, int Co, int KS, void* stream);
/* Dense(C<=32) + SparseCategoricalCrossentropy(from_logits) + accuracy count + dW/db/dh
 * (README.md:297, 301-302).  metrics += {loss sum, correct, rows}.  dh_colsum (optional, [K])
 * accumulates the bias gradient of the layer that produced h. */
int b2d_dense_xent_head(const float* h, const float* w, const float* bias, const int* labels,
                        float* logits_out, float* dh, float* dw, float* db, float* dh_colsum,
                        float* metrics, int B, int K, int C, float grad_scale, int relu_mask,
                        void* stream);
int b2d_softmax_xent(const float* logits, const int* labels, float* dlogits, float* metrics, int B,
                     int C, float grad_scale, void* stream);
int b2d_maxpool2_fwd(const float* x, float* y, uint8_t* code, int N, int H, int W, int C, void* stream);
int b2d_maxpool2_bwd(const float* dy, const uint8_t* code, const float* y_pooled, float* dx, int N,
                     int H, int W, int C, int act_relu, void* stream);
int b2d_colsum(const float* dy, float* db, int M, int N, void* stream);          /* db += column sums */
int b2d_relu_bwd(const float* dy, const float* y, float* dx, size_t n, void* stream);
/* Direct (CUDA-core) convolutions: any stride / zero padding.  ph, pw = padding at the top / left;
 * OH, OW = output size (<= 0: derive from symmetric padding) — TensorFlow's 'same' padding is
 * asymmetric for even totals, which the explicit output size expresses. */
int b2d_conv2d_fwd_direct(const float* x, const float* w, const float* bias, float* y, int N, int H,
                          int W, int Ci, int Co, int KH, int KW, int sh, int sw, int ph, int pw,
                          int OH, int OW, int relu, void* stream);
int b2d_conv2d_dgrad_direct(const float* dy, const float* w, float* dx, int N, int H, int W, int Ci,
                            int Co, int KH, int KW, int sh, int sw, int ph, int pw, int OH, int OW,
                            void* stream);
int b2d_conv2d_wgrad_direct(const float* x, const float* dy, float* dw, int N, int H, int W, int Ci,
                            int Co, int KH, int KW, int sh, int sw, int ph, int pw, int OH, int OW,
                            void* stream);
/* CUDA-core GEMM with arbitrary strides (tiny / unaligned problems). */
int b2d_sgemm_simt(const float* A, int64_t sam, int64_t sak, const float* B, int64_t sbk, int64_t sbn,
                   float* D, int64_t ldd, const float* bias, int M, int N, int K, int relu,
                   int accumulate, void* stream);
/* Single-replica optimizer step: w[:n_params] -= lr*g, w[n_params:] += g, g = 0 (README.md:302). */
int b2d_sgd_apply(float* w, float* g, size_t n_params, size_t n_total, float lr, void* stream);
int b2d_cast_f32_bf16(const float* src, void* dst, size_t n, void* stream);
/* ---------------------------------------------------------------------------------------
 * ResNet-family kernels (csrc/cuda/resnet_kernels.cu; BASELINE.json config 4 — none of these
 * layers has a counterpart in the reference README).  `dtype` is the storage type of the
 * activations (B2D_F32 or B2D_BF16, arithmetic is fp32); per-channel vectors stay fp32.
 * Channel counts must be multiples of the 16-byte vector: 4 (fp32) or 8 (bf16).
 * ------------------------------------------------------------------------------------- */
typedef struct b2d_bn_desc {
  int dtype;
  int M, C;                  /* x, y, dy, dx are [M rows][C channels]                              */
  int relu;                  /* fuse ReLU: y = max(.., 0); backward masks dy with y > 0             */
  int inference;             /* 1: normalise with moving_mean / moving_var, touch nothing else      */
  float eps, momentum;
  float inv_lr;              /* see moving_*_grad                                                   */
  const void* x;
  void* y;
  const void* dy;
  void* dx;                  /* NULL: parameter gradients only                                      */
  const float* gamma;
  const float* beta;
  /* Moving statistics [C] each.  Training with moving_*_grad == NULL updates them in place
   * (m = momentum*m + (1-momentum)*batch).  With moving_*_grad set, the kernel instead stores the
   * pseudo-gradient (1-momentum)*(m - batch)*inv_lr there, so that the optimizer's w -= lr*mean_r(g)
   * — fused into the cross-replica all-reduce — performs the same update with the batch statistics
   * AVERAGED OVER REPLICAS (what MirroredStrategy does for BatchNorm moving averages). */
  float* moving_mean;
  float* moving_var;
  float* moving_mean_grad;
  float* moving_var_grad;
  /* Scratch [2C] each, zero before the first call: `stats` (batch sums; cleared by backward),
   * `red` (backward sums; cleared by the next forward), `saved` (mean | rstd for backward). */
  float* stats;
  float* saved;
  float* red;
  float* dgamma;             /* += */
  float* dbeta;              /* += */
  /* Optional residual input (like x): forward computes y = act(bn(x) + residual) — the tail of a
   * residual block in one pass; backward then applies NO ReLU mask (the producer of dy, which is the
   * gradient of the block output, has already masked it). */
  const void* residual;
} b2d_bn_desc;
int b2d_bn_forward(const b2d_bn_desc* d, void* stream);    /* 2 launches (1 in inference mode)       */
int b2d_bn_backward(const b2d_bn_desc* d, void* stream);   /* 2 launches                             */
/* mode 0: out = a + b; 1: out = relu(a + b); 2: out = (a + b) where mask > 0 else 0.  b may be NULL. */
int b2d_add_act(int dtype, const void* a, const void* b, const void* mask, void* out, size_t n, int mode,
                void* stream);
/* General max pooling: window k, stride s, top/left padding p, explicit output size (<= 0: derive). */
int b2d_maxpool_fwd(int dtype, const void* x, void* y, uint8_t* code, int N, int H, int W, int C, int k, int s,
                    int p, int OH, int OW, void* stream);
int b2d_maxpool_bwd(int dtype, const void* dy, const uint8_t* code, void* dx, int N, int H, int W, int C, int k,
                    int s, int p, int OH, int OW, void* stream);
/* Global average pooling: y / dy are fp32 [N, C]; dx is zeroed where mask <= 0 (mask may be NULL). */
int b2d_gap_fwd(int dtype, const void* x, float* y, int N, int HW, int C, void* stream);
int b2d_gap_bwd(int dtype, const float* dy, const void* mask, void* dx, int N, int HW, int C, void* stream);
/* Patch matrix for small-channel convolutions (feeds the tcgen05 GEMM): col[pixel][kpad], kpad >= KH*KW*Ci. */
int b2d_im2col(int in_dtype, int out_dtype, const void* x, void* col, int N, int H, int W, int Ci, int KH, int KW,
               int sh, int sw, int ph, int pw, int OH, int OW, int kpad, void* stream);
int b2d_cast(int src_dtype, int dst_dtype, const void* src, void* dst, size_t n, void* stream);
int b2d_colsum_t(int dtype, const void* dy, float* db, int M, int C, void* stream);   /* db += column sums */

/* ---------------------------------------------------------------------------------------
 * Symmetric memory over NVLink (csrc/runtime/symm.cpp): one block per GPU, POSIX-fd exchange
 * over Unix sockets, every peer mapped into every process; optional NVLS multicast binding.
 * create / multicast_phase1 / multicast_phase2 are collective over the ranks of `tag`
 * (the caller provides a cross-rank barrier between the multicast phases).
 * ------------------------------------------------------------------------------------- */
typedef struct b2d_symm b2d_symm;
int b2d_symm_create(b2d_symm** out, const char* tag, int rank, int world, int device, size_t bytes,
                    double timeout_s);
int b2d_symm_multicast_phase1(b2d_symm* s);
int b2d_symm_multicast_phase2(b2d_symm* s);
void* b2d_symm_ptr(const b2d_symm* s, int rank);       /* this process' mapping of `rank`'s block */
void* b2d_symm_multicast_ptr(const b2d_symm* s);       /* NULL unless multicast is bound          */
size_t b2d_symm_bytes(const b2d_symm* s);
int b2d_symm_multicast_supported(const b2d_symm* s);
int b2d_symm_destroy(b2d_symm* s);

/* ---------------------------------------------------------------------------------------
 * Fused gradient all-reduce + 1/world averaging + SGD + metric sums + gradient zeroing over
 * peer memory (csrc/cuda/allreduce.cu): replaces "batch_all_reduce: 6 all-reduces", the four
 * metric all-reduces and the optimizer step of the reference (README.md:403-407, 302).
 * algo: 0 one-shot, 1 two-shot (weights all-gather), 2 NVLS multicast, 3 two-shot pull, -1 by size;
 * 4 = the gradients were already reduce-scattered by the producing GEMM (b2d_gemm_plan_set_scatter): every
 * owner updates its shard from its own arena and broadcasts the weights.
 * ------------------------------------------------------------------------------------- */
typedef struct b2d_allreduce_desc {
  void* w[8];      /* weight arenas of every rank (peer-mapped)          */
  void* g[8];      /* gradient arenas                                    */
  void* sig[8];    /* signal pads, b2d_allreduce_signal_bytes() each     */
  void* epoch[8];  /* barrier epochs, b2d_allreduce_epoch_bytes() each   */
  void* out[8];    /* plain one-shot all-reduce destinations             */
  void* w_mc;      /* multicast views (NVLS) or NULL                     */
  void* g_mc;
  int64_t n_params; /* floats updated by SGD (multiple of 4)             */
  int64_t n_total;  /* n_params + metric tail slots (multiple of 4)      */
  float lr;
  int rank, world, algo;
  int emulate;      /* all ranks in one cooperative launch on one GPU (tests) */
  int plain;        /* sum only: no SGD, no zeroing                     */
  int blocks;       /* 0 = automatic                                    */
  void* g_zero;     /* double-buffered gradients: this rank's OTHER gradient buffer, zeroed by
                       the kernel (one-shot then runs with a single cross-GPU barrier); NULL =
                       single buffer, zeroed in place after a second barrier              */
  void* timeline;   /* optional debug, 8 x int64 per block or NULL: launches, ns waited for upstream / in the
                       release / for peers / moving data (sums), then start, past-barrier and end stamps of
                       the last launch (globaltimer)                                            */
  int64_t seg_begin[2]; /* bucketing: up to two [begin, end) float ranges of the arenas handled  */
  int64_t seg_end[2];   /* by this launch (multiples of 4); nseg == 0 means the whole arena      */
  int nseg;
  void* w_out[8];   /* weight ping-pong: where the UPDATED weights are written (every rank's other
                       weight buffer); NULL entries = update w in place                       */
  void* w_out_mc;   /* multicast view of w_out (NVLS) or NULL                                 */
  int flags;        /* B2D_AR_* bits                                                          */
} b2d_allreduce_desc;
#define B2D_AR_NO_EXIT_BARRIER 1 /* the caller closes the step with a later exchange (b2d_ll_allreduce_sgd):
                                    skip the barrier after the remote weight writes; needs g_zero    */
#define B2D_AR_STRICT_ENTRY 2    /* entry barrier with st.release.sys (also: env B200DIST_AR_STRICT=1) */
size_t b2d_allreduce_signal_bytes(void);
size_t b2d_allreduce_epoch_bytes(void);
int b2d_allreduce_pick_algo(int64_t n_total_floats, int world, int have_multicast);
int b2d_allreduce_sgd(const b2d_allreduce_desc* desc, void* stream);

/* Small-bucket variant (the first layer's gradients + the metric sums, README.md:404-407): every
 * rank pushes {value, epoch} pairs as single 8-byte stores into the peers' inboxes and sums what
 * arrives in rank order — one one-way NVLink trip, no barrier, bit-identical on every rank; the
 * exchange doubles as the closing barrier of a step whose other buckets used
 * B2D_AR_NO_EXIT_BARRIER.  At most b2d_ll_max_floats() values per launch. */
typedef struct b2d_ll_desc {
  void* inbox[8];   /* b2d_ll_inbox_bytes() of zero-initialised symmetric memory per rank        */
  void* epoch[8];   /* b2d_ll_epoch_bytes() of zero-initialised local memory per rank           */
  void* w[8];       /* weights read  (only [rank] unless emulate)                               */
  void* w_out[8];   /* weights written (NULL = in place)                                        */
  void* g[8];       /* gradients of this step                                                   */
  void* g_zero[8];  /* the other gradient buffer, zeroed over the same ranges, or NULL          */
  int64_t seg_begin[2], seg_end[2]; /* one or two [begin, end) float ranges of the arenas       */
  int nseg;
  int64_t n_params; /* indices >= n_params are metric sums: added, not averaged                 */
  float lr;
  int rank, world;
  int emulate;      /* all ranks in one cooperative launch on one GPU (tests)                   */
  void* timeline;   /* optional debug: 4 x int64 globaltimer stamps per block (start, inputs ready, done) */
} b2d_ll_desc;
size_t b2d_ll_inbox_bytes(void);
size_t b2d_ll_epoch_bytes(void);
int64_t b2d_ll_max_floats(void);
int b2d_ll_allreduce_sgd(const b2d_ll_desc* desc, void* stream);

/* ---------------------------------------------------------------------------------------
 * Native step executor (csrc/runtime/executor.cpp): captures the step's launches into CUDA
 * graphs (one per input staging slot) and runs the fit() hot loop — input prefetch on a copy
 * stream, graph launch, per-step device->host read of the metric sums (README.md:304, 392).
 * ------------------------------------------------------------------------------------- */
typedef struct b2d_executor b2d_executor;
int b2d_exec_create(b2d_executor** out);
void b2d_exec_destroy(b2d_executor* e);
void* b2d_exec_stream(b2d_executor* e);
void* b2d_exec_side_stream(b2d_executor* e);   /* concurrent branch of a step (bucketed all-reduce) */
int b2d_exec_fork(b2d_executor* e);            /* side stream starts after everything on main so far */
int b2d_exec_join(b2d_executor* e);            /* main continues after everything on side so far     */
int b2d_exec_set_io(b2d_executor* e, int slot, void* x_dev, size_t x_bytes, void* y_dev, size_t y_bytes);
int b2d_exec_set_metrics(b2d_executor* e, const void* metrics_dev, size_t bytes, void* pinned_ring,
                         int ring_len);
int b2d_exec_set_metrics_slot(b2d_executor* e, int slot, const void* metrics_dev); /* weight ping-pong */
int b2d_exec_begin_capture(b2d_executor* e);
int b2d_exec_end_capture(b2d_executor* e, int slot);
uint64_t b2d_exec_graph_nodes(const b2d_executor* e, int slot);
int b2d_exec_run(b2d_executor* e, const void* x_src, size_t x_stride, const void* y_src, size_t y_stride,
                 int64_t first_step, int64_t steps_in_src, int64_t n_steps, int sync, float* elapsed_ms);
int b2d_exec_sync(b2d_executor* e);
uint64_t b2d_exec_steps_done(const b2d_executor* e);
int b2d_host_alloc(void** out, size_t bytes);   /* pinned host memory */
int b2d_host_free(void* p);

/* ---------------------------------------------------------------------------------------
 * CPU backend (libb200dist_cpu.so): C++ reference ops + shared-memory fused all-reduce/SGD.
 * ------------------------------------------------------------------------------------- */
void b2c_set_num_threads(int n);
/* co-located worker processes: `threads` pool threads pinned to CPUs first_cpu.. of the affinity mask */
void b2c_set_thread_layout(int threads, int first_cpu);
/* ph, pw = padding at the top / left; OH, OW = output size (<= 0: derive from symmetric padding). */
void b2c_conv2d_fwd(const float* x, const float* w, const float* bias, float* y, int N, int H, int W,
                    int Ci, int Co, int KH, int KW, int sh, int sw, int ph, int pw, int OH, int OW, int relu);
void b2c_conv2d_wgrad(const float* x, const float* dy, float* dw, float* db, int N, int H, int W, int Ci,
                      int Co, int KH, int KW, int sh, int sw, int ph, int pw, int OH, int OW);
void b2c_conv2d_dgrad(const float* dy, const float* w, float* dx, int N, int H, int W, int Ci, int Co,
                      int KH, int KW, int sh, int sw, int ph, int pw, int OH, int OW);
void b2c_maxpool2_fwd(const float* x, float* y, uint8_t* code, int N, int H, int W, int C);
void b2c_maxpool2_bwd(const float* dy, const uint8_t* code, const float* y_pooled, float* dx, int N,
                      int H, int W, int C, int act_relu);
void b2c_gemm(const float* A, int64_t sam, int64_t sak, const float* B, int64_t sbk, int64_t sbn, float* D,
              int64_t ldd, const float* bias, int M, int N, int K, int relu, int accumulate);
void b2c_softmax_xent(const float* z, const int32_t* labels, float* dz, float* metrics, int B, int C,
                      float grad_scale);
void b2c_relu_bwd(const float* dy, const float* y, float* dx, int64_t n);
void b2c_colsum(const float* dy, float* db, int M, int N);
void b2c_sgd_apply(float* w, float* g, int64_t n_params, int64_t n_total, float lr);
/* ResNet-family layers on the CPU (csrc/cpu/resnet_ops.cpp): same semantics as the b2d_* kernels above. */
void b2c_bn_fwd(const float* x, const float* gamma, const float* beta, float* y, float* saved, float* moving_mean,
                float* moving_var, float* g_mean, float* g_var, int64_t M, int C, float eps, float momentum,
                float inv_lr, int relu, int inference, const float* residual);
void b2c_bn_bwd(const float* x, const float* y, const float* dy, const float* gamma, const float* saved, float* dx,
                float* dgamma, float* dbeta, int64_t M, int C, int relu);
void b2c_add_act(const float* a, const float* b, const float* mask, float* out, int64_t n, int mode);
void b2c_maxpool_fwd(const float* x, float* y, uint8_t* code, int N, int H, int W, int C, int k, int s, int p,
                     int OH, int OW);
void b2c_maxpool_bwd(const float* dy, const uint8_t* code, float* dx, int N, int H, int W, int C, int k, int s,
                     int p, int OH, int OW);
void b2c_gap_fwd(const float* x, float* y, int N, int HW, int C);
void b2c_gap_bwd(const float* dy, const float* mask, float* dx, int N, int HW, int C);
typedef struct b2c_shm_comm b2c_shm_comm;
const char* b2c_last_error(void);
b2c_shm_comm* b2c_shm_create(const char* name, int rank, int world, int64_t capacity_floats, double timeout_s);
void b2c_shm_destroy(b2c_shm_comm* c);
int b2c_shm_barrier(b2c_shm_comm* c);
int b2c_shm_allreduce(b2c_shm_comm* c, float* buf, int64_t n);
int b2c_shm_allreduce_sgd(b2c_shm_comm* c, float* w, float* g, int64_t n_params, int64_t n_total, float lr);

#ifdef __cplusplus
}
#endif
#endif /* B200DIST_H_ */
