"""tcgen05 GEMM plans over torch CUDA tensors (csrc/cuda/gemm_tcgen05.cu).

A plan freezes pointers, shapes and TMA descriptors, so running it is a single kernel
launch with no host-side work — which is what makes the whole training step capturable in
one CUDA graph.  Covers Dense forward / dgrad / wgrad (README.md:296-297).
"""
from __future__ import annotations

import ctypes as C

import torch

from .._native import cuda_lib as nl

F32, BF16 = 0, 1
_DT = {torch.float32: F32, torch.bfloat16: BF16}


def _stream_ptr(stream=None) -> int:
    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream


class GemmPlan:
    """D[M,N] = epilogue(A[M,K] @ B[K,N]).

    ``a`` is either ``[M, K]`` (K-major, ``a_mn_major=False``) or ``[K, M]`` (MN-major);
    ``b`` is either ``[N, K]`` (K-major) or ``[K, N]`` (MN-major, the Keras Dense kernel layout).
    Tensors must be row-contiguous in their last dimension; the row pitch is taken from
    ``stride(0)``.
    """

    def __init__(self, a, b, d, *, a_mn_major=False, b_mn_major=False, bias=None, mask=None,
                 relu=False, accumulate=False, splits=0):
        assert a.is_cuda and b.is_cuda and d.is_cuda
        assert a.dtype == b.dtype and a.dtype in _DT and d.dtype in _DT
        assert a.stride(1) == 1 and b.stride(1) == 1 and d.stride(1) == 1
        if a_mn_major:
            K, M = a.shape
        else:
            M, K = a.shape
        if b_mn_major:
            Kb, N = b.shape
        else:
            N, Kb = b.shape
        assert K == Kb, f"K mismatch: {K} vs {Kb}"
        assert tuple(d.shape) == (M, N), f"D shape {tuple(d.shape)} != {(M, N)}"
        desc = nl.GemmDesc()
        desc.dtype, desc.out_dtype = _DT[a.dtype], _DT[d.dtype]
        desc.M, desc.N, desc.K = M, N, K
        desc.A, desc.lda, desc.a_mn_major = a.data_ptr(), a.stride(0), int(a_mn_major)
        desc.B, desc.ldb, desc.b_mn_major = b.data_ptr(), b.stride(0), int(b_mn_major)
        desc.D, desc.ldd = d.data_ptr(), d.stride(0)
        desc.bias = bias.data_ptr() if bias is not None else None
        if mask is not None:
            assert mask.dtype in _DT and mask.stride(1) == 1
            desc.mask, desc.ldmask, desc.mask_dtype = mask.data_ptr(), mask.stride(0), _DT[mask.dtype]
        desc.relu, desc.accumulate, desc.splits = int(relu), int(accumulate), int(splits)
        lib = nl.lib()
        self.geometry = nl.GemmGeometry()
        nl.check(lib.b2d_gemm_query(C.byref(desc), C.byref(self.geometry)), "b2d_gemm_query")
        ws_bytes = lib.b2d_gemm_workspace_bytes(C.byref(desc))
        self._keep = [a, b, d, bias, mask]
        if ws_bytes:
            # Scratch for the split-K rendezvous; zeroed once, left zero by every launch.
            self.workspace = torch.zeros(ws_bytes // 4, dtype=torch.float32, device=a.device)
            self.counters = torch.zeros(lib.b2d_gemm_counter_count(C.byref(desc)),
                                        dtype=torch.int32, device=a.device)
            desc.workspace, desc.counters = self.workspace.data_ptr(), self.counters.data_ptr()
        else:
            self.workspace = self.counters = None
        self._desc = desc
        h = C.c_void_p()
        nl.check(lib.b2d_gemm_plan_create(C.byref(desc), C.byref(h)), "b2d_gemm_plan_create")
        self._h = h
        self._lib = lib

    def set_scatter(self, bases, d_off: int, seg_begin: int, per: int, rank: int) -> None:
        """GEMM -> reduce-scatter: add every 16-byte vector of the result into the gradient arena of the
        replica that owns it (``bases``: every replica's arena as mapped here; offsets in floats)."""
        sc = nl.GemmScatter()
        for p, b in enumerate(bases):
            sc.base[p] = b
        sc.d_off, sc.seg_begin, sc.per, sc.rank, sc.world = d_off, seg_begin, per, rank, len(bases)
        nl.check(self._lib.b2d_gemm_plan_set_scatter(self._h, C.byref(sc)), "b2d_gemm_plan_set_scatter")

    def run(self, stream=None) -> None:
        nl.check(self._lib.b2d_gemm_plan_run(self._h, _stream_ptr(stream)), "b2d_gemm_plan_run")

    __call__ = run

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            self._lib.b2d_gemm_plan_destroy(h)
