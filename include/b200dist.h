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
  const float* mask;  /* optional, fp32 [M][ldmask]                        */
  int64_t ldmask;
  int relu;
  int accumulate;     /* 1: D += result (fp32 atomics; D must be fp32)     */
  int splits;         /* 0 = choose automatically                          */
  float* workspace;   /* split-K scratch, b2d_gemm_workspace_bytes()       */
  uint32_t* counters; /* split-K tile counters, b2d_gemm_counter_count()   */
} b2d_gemm_desc;

typedef struct b2d_gemm_plan b2d_gemm_plan;

/* Resolved launch geometry for a descriptor (splits after clamping etc.). */
typedef struct b2d_gemm_geometry {
  int block_n, m_tiles, n_tiles, splits, k_blocks, stages, smem_bytes;
} b2d_gemm_geometry;

int b2d_gemm_query(const b2d_gemm_desc* desc, b2d_gemm_geometry* out);
size_t b2d_gemm_workspace_bytes(const b2d_gemm_desc* desc);
size_t b2d_gemm_counter_count(const b2d_gemm_desc* desc);
/* Encodes the TMA descriptors (host only, no device allocation). */
int b2d_gemm_plan_create(const b2d_gemm_desc* desc, b2d_gemm_plan** out);
int b2d_gemm_plan_run(const b2d_gemm_plan* plan, void* stream);
void b2d_gemm_plan_destroy(b2d_gemm_plan* plan);

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

#ifdef __cplusplus
}
#endif
#endif /* B200DIST_H_ */
