"""The allocation-free GEMM entry point, driven through the C ABI exactly as a C host would.

Who needs it: programs that cannot allocate device memory where they call the GEMM — the GPU
belongs to another pool, or the call sits where a driver allocation is forbidden or too slow.
``b2d_gemm_noalloc`` keeps the split-K scratch in a fixed array inside the library, so the
caller passes only inputs and the output.  This test checks it against the regular entry
point (``b2d_gemm_plan_create`` / ``b2d_gemm_plan_run`` with a caller-provided workspace),
on back-to-back calls that reuse the shared array, including a change of shape in between,
and checks that an oversized problem is refused rather than overrunning the array.

Collected last on purpose (file name): a run that stops at the first failure has then already
exercised every other test.  It has no dependency on them and passes in a fresh process.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 2e-3      # same relative tolerance as tests/test_gemm_gpu.py (tf32 operands)


def _lib():
    from distributed_b200._native import cuda_lib as nl
    lib = nl.lib()
    lib.b2d_gemm_noalloc.argtypes = [C.POINTER(nl.GemmDesc), C.c_void_p]
    lib.b2d_gemm_noalloc.restype = C.c_int
    lib.b2d_gemm_noalloc_max_workspace_floats.restype = C.c_size_t
    lib.b2d_gemm_noalloc_max_tiles.restype = C.c_size_t
    return nl, lib


def _desc(nl, a, b, d, bias, relu, splits=0):
    """Dense-forward shaped problem: D[M,N] = relu(A[M,K] . B[K,N] + bias), B is (in, out)."""
    desc = nl.GemmDesc()
    desc.dtype, desc.out_dtype = 0, 0
    desc.M, desc.K = a.shape
    desc.N = b.shape[1]
    desc.A, desc.lda, desc.a_mn_major = a.data_ptr(), a.stride(0), 0
    desc.B, desc.ldb, desc.b_mn_major = b.data_ptr(), b.stride(0), 1
    desc.D, desc.ldd = d.data_ptr(), d.stride(0)
    desc.bias = bias.data_ptr()
    desc.relu, desc.accumulate, desc.splits = int(relu), 0, splits
    return desc


def _problem(M, N, K, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    a = torch.randn((M, K), device="cuda", generator=g) * 0.5
    b = torch.randn((K, N), device="cuda", generator=g) * 0.5
    bias = torch.randn((N,), device="cuda", generator=g)
    return a, b, bias


def _run_both(nl, lib, M, N, K, seed, check_driver_memory=False):
    from distributed_b200.ops.gemm import GemmPlan
    a, b, bias = _problem(M, N, K, seed)
    expected = torch.empty((M, N), device="cuda")
    plan = GemmPlan(a, b, expected, b_mn_major=True, bias=bias, relu=True)      # existing entry point
    assert plan.geometry.splits > 1, "the case must exercise the split-K scratch"
    plan.run()
    got = torch.full((M, N), float("nan"), device="cuda")                       # output owned by the caller
    stream = torch.cuda.current_stream().cuda_stream
    desc = _desc(nl, a, b, got, bias, True)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]                                  # driver-level free bytes
    rc = lib.b2d_gemm_noalloc(C.byref(desc), stream)
    assert rc == 0, lib.b2d_last_error().decode()
    torch.cuda.synchronize()
    if check_driver_memory:     # (not on the very first call: kernels may still be loading lazily)
        assert torch.cuda.mem_get_info()[0] == free_before, "the entry point changed the free device memory"
    scale = expected.abs().max().item()
    err = (got - expected).abs().max().item()
    assert err <= TOL * scale, f"M={M} N={N} K={K}: max abs diff {err} vs scale {scale}"
    # and both agree with an fp32 reference of the op
    ref = torch.relu(a.double() @ b.double() + bias.double()).float()
    assert (got - ref).abs().max().item() <= TOL * scale


def test_noalloc_gemm_matches_regular_entry_point_on_consecutive_calls():
    nl, lib = _lib()
    _run_both(nl, lib, 64, 64, 5408, seed=11)       # README dense-64 forward
    _run_both(nl, lib, 64, 64, 5408, seed=12, check_driver_memory=True)   # straight again: shared array reused
    _run_both(nl, lib, 200, 128, 4096, seed=13)     # other shape: stale contents must not leak in
    _run_both(nl, lib, 64, 64, 5408, seed=14, check_driver_memory=True)


def test_noalloc_gemm_rejects_problems_beyond_the_builtin_scratch():
    nl, lib = _lib()
    max_floats = lib.b2d_gemm_noalloc_max_workspace_floats()
    assert max_floats == 256 * 1024 and lib.b2d_gemm_noalloc_max_tiles() == 256
    M, N, K = 1024, 1024, 8192                       # forced split-K: 1 Mi partial sums > 256 Ki
    a, b, bias = _problem(M, N, K, seed=15)
    d = torch.zeros((M, N), device="cuda")
    rc = lib.b2d_gemm_noalloc(C.byref(_desc(nl, a, b, d, bias, False, splits=2)),
                              torch.cuda.current_stream().cuda_stream)
    assert rc == -4                                   # B2D_ERR_LIMIT
    msg = lib.b2d_last_error().decode()
    assert "exceeds the built-in maximum" in msg and "b2d_gemm_plan_create" in msg
    torch.cuda.synchronize()
    assert d.abs().max().item() == 0                  # nothing was launched
