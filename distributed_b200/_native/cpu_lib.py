"""ctypes bindings for ``libb200dist_cpu.so`` (C++ CPU ops + shared-memory all-reduce)."""
from __future__ import annotations

import ctypes as C
from pathlib import Path

LIB_PATH = Path(__file__).resolve().parents[1] / "_lib" / "libb200dist_cpu.so"
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            from . import build
            build.build_cpu()
        l = C.CDLL(str(LIB_PATH))
        vp, i, f, i64, d = C.c_void_p, C.c_int, C.c_float, C.c_int64, C.c_double
        conv = [i] * 13
        l.b2c_set_num_threads.argtypes = [i]
        l.b2c_set_thread_layout.argtypes = [i, i]
        l.b2c_set_thread_layout.restype = None
        l.b2c_conv2d_fwd.argtypes = [vp, vp, vp, vp, *conv, i]
        l.b2c_conv2d_wgrad.argtypes = [vp, vp, vp, vp, *conv]
        l.b2c_conv2d_dgrad.argtypes = [vp, vp, vp, *conv]
        l.b2c_maxpool2_fwd.argtypes = [vp, vp, vp, i, i, i, i]
        l.b2c_maxpool2_bwd.argtypes = [vp, vp, vp, vp, i, i, i, i, i]
        l.b2c_gemm.argtypes = [vp, i64, i64, vp, i64, i64, vp, i64, vp, i, i, i, i, i]
        l.b2c_softmax_xent.argtypes = [vp, vp, vp, vp, i, i, f]
        l.b2c_relu_bwd.argtypes = [vp, vp, vp, i64]
        l.b2c_colsum.argtypes = [vp, vp, i, i]
        l.b2c_sgd_apply.argtypes = [vp, vp, i64, i64, f]
        l.b2c_bn_fwd.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp, i64, i, f, f, f, i, i, vp]
        l.b2c_bn_bwd.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, i64, i, i]
        l.b2c_add_act.argtypes = [vp, vp, vp, vp, i64, i]
        l.b2c_maxpool_fwd.argtypes = [vp, vp, vp, i, i, i, i, i, i, i, i, i]
        l.b2c_maxpool_bwd.argtypes = [vp, vp, vp, i, i, i, i, i, i, i, i, i]
        l.b2c_gap_fwd.argtypes = [vp, vp, i, i, i]
        l.b2c_gap_bwd.argtypes = [vp, vp, vp, i, i, i]
        for fn in ("b2c_conv2d_fwd", "b2c_conv2d_wgrad", "b2c_conv2d_dgrad", "b2c_maxpool2_fwd",
                   "b2c_maxpool2_bwd", "b2c_gemm", "b2c_softmax_xent", "b2c_relu_bwd", "b2c_colsum",
                   "b2c_sgd_apply", "b2c_set_num_threads", "b2c_bn_fwd", "b2c_bn_bwd", "b2c_add_act",
                   "b2c_maxpool_fwd", "b2c_maxpool_bwd", "b2c_gap_fwd", "b2c_gap_bwd"):
            getattr(l, fn).restype = None
        l.b2c_last_error.restype = C.c_char_p
        l.b2c_shm_create.argtypes = [C.c_char_p, i, i, i64, d]
        l.b2c_shm_create.restype = vp
        l.b2c_shm_destroy.argtypes = [vp]
        l.b2c_shm_destroy.restype = None
        l.b2c_shm_barrier.argtypes = [vp]
        l.b2c_shm_allreduce.argtypes = [vp, vp, i64]
        l.b2c_shm_allreduce_sgd.argtypes = [vp, vp, vp, i64, i64, f]
        _lib = l
    return _lib
