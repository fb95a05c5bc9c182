"""Implicit-GEMM convolution plans on the tcgen05 core (csrc/cuda/gemm_tcgen05.cu).

Stride-1 NHWC x HWIO convolutions: the A operand is the activation tensor itself, read through
4-D TMA boxes shifted per filter tap (zero padding = TMA out-of-bounds fill), so there is no
im2col buffer.  Forward, data gradient and (ConvWgradPlan) filter gradient all run on the
tensor cores.  Conv2D of the README / scaled models (README.md:293, BASELINE.json config 3).
"""
from __future__ import annotations

import ctypes as C

import torch

from .._native import cuda_lib as nl
from .gemm import _DT, _stream_ptr


class ConvDesc(C.Structure):
    _fields_ = [("dtype", C.c_int), ("out_dtype", C.c_int),
                ("N", C.c_int), ("H", C.c_int), ("W", C.c_int), ("Ci", C.c_int), ("Co", C.c_int),
                ("KH", C.c_int), ("KW", C.c_int), ("pad_h", C.c_int), ("pad_w", C.c_int),
                ("x", C.c_void_p), ("w", C.c_void_p), ("y", C.c_void_p),
                ("bias", C.c_void_p), ("mask", C.c_void_p), ("relu", C.c_int),
                ("stride", C.c_int), ("OH", C.c_int), ("OW", C.c_int), ("mask_dtype", C.c_int),
                ("addend", C.c_void_p)]


_declared = False


def _lib():
    global _declared
    lib = nl.lib()
    if not _declared:
        lib.b2d_conv_tc_supported.argtypes = [C.POINTER(ConvDesc), C.c_int]
        lib.b2d_conv_fwd_plan_create.argtypes = [C.POINTER(ConvDesc), C.POINTER(C.c_void_p)]
        lib.b2d_conv_dgrad_plan_create.argtypes = [C.POINTER(ConvDesc), C.POINTER(C.c_void_p)]
        lib.b2d_conv_dgrad_class_plan_create.argtypes = [C.POINTER(ConvDesc), C.c_int, C.c_int, C.POINTER(C.c_void_p)]
        lib.b2d_memset_async.argtypes = [C.c_void_p, C.c_int, C.c_size_t, C.c_void_p]
        lib.b2d_conv_wgrad_plan_create.argtypes = [C.POINTER(ConvDesc), C.POINTER(C.c_void_p)]
        lib.b2d_conv_wgrad_plan_run.argtypes = [C.c_void_p, C.c_void_p]
        lib.b2d_conv_wgrad_plan_destroy.argtypes = [C.c_void_p]
        lib.b2d_conv_wgrad_plan_destroy.restype = None
        _declared = True
    return lib


def _desc(x, w, y, padding, bias=None, mask=None, relu=False, stride=1, addend=None):
    N, H, W, Ci = x.shape
    KH, KW, Ci2, Co = w.shape
    assert Ci == Ci2
    d = ConvDesc()
    d.dtype, d.out_dtype = _DT[x.dtype], _DT[y.dtype]
    d.N, d.H, d.W, d.Ci, d.Co, d.KH, d.KW = N, H, W, Ci, Co, KH, KW
    d.pad_h, d.pad_w = padding
    d.x, d.w, d.y = x.data_ptr(), w.data_ptr(), y.data_ptr()
    d.bias = bias.data_ptr() if bias is not None else None
    d.mask = mask.data_ptr() if mask is not None else None
    d.mask_dtype = _DT[mask.dtype] if mask is not None else 0
    d.relu = int(relu)
    d.addend = addend.data_ptr() if addend is not None else None
    d.stride, d.OH, d.OW = int(stride), y.shape[1], y.shape[2]
    return d


def tc_supported(x_shape, w_shape, padding, dtype, dgrad=False, strides=(1, 1), out_hw=None) -> bool:
    if strides[0] != strides[1] or dtype not in _DT:
        return False
    d = ConvDesc()
    d.stride = int(strides[0])
    if out_hw is not None:
        d.OH, d.OW = out_hw
    d.dtype = d.out_dtype = _DT[dtype]
    d.N, d.H, d.W, d.Ci = x_shape
    d.KH, d.KW, _, d.Co = w_shape
    d.pad_h, d.pad_w = padding
    return bool(_lib().b2d_conv_tc_supported(C.byref(d), int(dgrad)))


class _Plan:
    def __init__(self, create, desc, keep):
        self._lib = _lib()
        h = C.c_void_p()
        nl.check(create(C.byref(desc), C.byref(h)), "conv plan create")
        self._h, self._keep = h, keep

    def run(self, stream=None):
        nl.check(self._lib.b2d_gemm_plan_run(self._h, _stream_ptr(stream)), "conv plan run")

    __call__ = run

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            self._lib.b2d_gemm_plan_destroy(h)


class ConvFwdPlan(_Plan):
    """y = act(conv2d(x, w) + bias), x [N,H,W,Ci], w [KH,KW,Ci,Co], y [N,OH,OW,Co]."""

    def __init__(self, x, w, y, padding=(0, 0), bias=None, relu=False, stride=1):
        super().__init__(_lib().b2d_conv_fwd_plan_create, _desc(x, w, y, padding, bias, None, relu, stride),
                         [x, w, y, bias])


class ConvDgradPlan(_Plan):
    """dx = conv2d_transpose(dy, w) [+ addend] (zeroed where mask <= 0), dx [N,H,W,Ci], dy [N,OH,OW,Co].
    ``addend`` (same shape / dtype as dx) is summed in the epilogue: the other branch's gradient of a
    residual block, so that the block's gradient sum costs no extra pass."""

    def __init__(self, dy, w, dx, padding=(0, 0), mask=None, addend=None):
        assert addend is None or (addend.shape == dx.shape and addend.dtype == dx.dtype)
        super().__init__(_lib().b2d_conv_dgrad_plan_create,
                         _desc(dx, w, dy, padding, None, mask, False, addend=addend), [dy, w, dx, mask, addend])


class ConvDgradStridedPlan:
    """dx for a strided convolution: stride^2 unit-stride sub-convolutions of dy, one per parity
    class of the output pixel; classes without taps are covered by clearing dx first."""

    def __init__(self, dy, w, dx, padding, stride, mask=None):
        self._lib = _lib()
        self._keep = [dy, w, dx, mask]
        d = _desc(dx, w, dy, padding, None, mask, False, stride)
        self._plans, self._need_clear = [], False
        for a in range(stride):
            for b in range(stride):
                h = C.c_void_p()
                nl.check(self._lib.b2d_conv_dgrad_class_plan_create(C.byref(d), a, b, C.byref(h)),
                         "conv dgrad class plan create")
                if h.value:
                    self._plans.append(h)
                else:
                    self._need_clear = True
        self._dx = dx

    def run(self, stream=None):
        st = _stream_ptr(stream)
        if self._need_clear:
            nl.check(self._lib.b2d_memset_async(self._dx.data_ptr(), 0, self._dx.numel() * self._dx.element_size(), st),
                     "memset")
        for h in self._plans:
            nl.check(self._lib.b2d_gemm_plan_run(h, st), "conv dgrad class run")

    __call__ = run

    def __del__(self):
        for h in getattr(self, "_plans", []):
            self._lib.b2d_gemm_plan_destroy(h)
        self._plans = []


def dgrad_strided_supported(dx_shape, w_shape, stride, dtype) -> bool:
    if dtype not in _DT or stride < 2:
        return False
    es = 4 if dtype == torch.float32 else 2
    _, H, W, Ci = dx_shape
    Co = w_shape[3]
    return (Co * es) % 128 == 0 and Ci % 4 == 0 and (W + stride - 1) // stride <= 128


class ConvWgradPlan:
    """dw += x (*) dy over all pixels (fp32 accumulation into dw [KH,KW,Ci,Co])."""

    def __init__(self, x, dy, dw, padding=(0, 0), stride=1):
        self._lib = _lib()
        d = _desc(x, dw, dy, padding, stride=stride)
        d.dtype = _DT[x.dtype]
        h = C.c_void_p()
        nl.check(self._lib.b2d_conv_wgrad_plan_create(C.byref(d), C.byref(h)), "conv wgrad plan create")
        self._h, self._keep = h, [x, dy, dw]

    def run(self, stream=None):
        nl.check(self._lib.b2d_conv_wgrad_plan_run(self._h, _stream_ptr(stream)), "conv wgrad run")

    __call__ = run

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            self._lib.b2d_conv_wgrad_plan_destroy(h)
