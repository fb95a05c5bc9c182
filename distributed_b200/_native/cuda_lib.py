"""ctypes bindings for ``libb200dist_cuda.so`` (C ABI declared in ``include/b200dist.h``).

The library is loaded lazily.  On a machine with a GPU a missing library is a hard error —
the CUDA backend never falls back to PyTorch kernels silently.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import os

# B200DIST_CUDA_LIB: load another build of the library (A/B measurements of two builds on one box)
LIB_PATH = Path(os.environ.get("B200DIST_CUDA_LIB") or
                Path(__file__).resolve().parents[1] / "_lib" / "libb200dist_cuda.so")

_lib = None


class NativeError(RuntimeError):
    pass


class GemmDesc(C.Structure):
    _fields_ = [
        ("dtype", C.c_int), ("out_dtype", C.c_int),
        ("M", C.c_int), ("N", C.c_int), ("K", C.c_int),
        ("A", C.c_void_p), ("lda", C.c_int64), ("a_mn_major", C.c_int),
        ("B", C.c_void_p), ("ldb", C.c_int64), ("b_mn_major", C.c_int),
        ("D", C.c_void_p), ("ldd", C.c_int64),
        ("bias", C.c_void_p), ("mask", C.c_void_p), ("ldmask", C.c_int64),
        ("relu", C.c_int), ("accumulate", C.c_int), ("splits", C.c_int),
        ("workspace", C.c_void_p), ("counters", C.c_void_p),
        ("mask_dtype", C.c_int),
    ]


class GemmScatter(C.Structure):
    _fields_ = [("base", C.c_void_p * 8), ("d_off", C.c_int64), ("seg_begin", C.c_int64), ("per", C.c_int64),
                ("rank", C.c_int), ("world", C.c_int)]


class GemmGeometry(C.Structure):
    _fields_ = [(n, C.c_int) for n in
                ("block_n", "m_tiles", "n_tiles", "splits", "k_blocks", "stages", "smem_bytes", "ctas", "pair")]


def _declare(lib) -> None:
    lib.b2d_last_error.restype = C.c_char_p
    lib.b2d_launch_count.restype = C.c_uint64
    lib.b2d_gemm_query.argtypes = [C.POINTER(GemmDesc), C.POINTER(GemmGeometry)]
    lib.b2d_gemm_workspace_bytes.argtypes = [C.POINTER(GemmDesc)]
    lib.b2d_gemm_workspace_bytes.restype = C.c_size_t
    lib.b2d_gemm_counter_count.argtypes = [C.POINTER(GemmDesc)]
    lib.b2d_gemm_counter_count.restype = C.c_size_t
    lib.b2d_gemm_plan_create.argtypes = [C.POINTER(GemmDesc), C.POINTER(C.c_void_p)]
    lib.b2d_gemm_plan_run.argtypes = [C.c_void_p, C.c_void_p]
    lib.b2d_gemm_plan_destroy.argtypes = [C.c_void_p]
    lib.b2d_gemm_plan_destroy.restype = None
    lib.b2d_gemm_plan_set_scatter.argtypes = [C.c_void_p, C.POINTER(GemmScatter)]
    # Everything else is declared by the op wrappers next to their use (see ops/).


def lib():
    """Return the loaded CUDA library, building nothing; raise if it is missing."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise NativeError(
                f"{LIB_PATH} is missing: run `python -m distributed_b200._native.build` "
                "(or __graft_entry__.build()) before using the CUDA backend")
        _lib = C.CDLL(str(LIB_PATH), mode=C.RTLD_GLOBAL)
        _declare(_lib)
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().b2d_last_error().decode(errors="replace")
        raise NativeError(f"{what or 'native call'} failed (rc={rc}): {msg}")


def launch_count() -> int:
    return int(lib().b2d_launch_count())
