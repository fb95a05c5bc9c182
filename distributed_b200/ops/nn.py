"""ctypes wrappers for the non-GEMM sm_100a kernels (csrc/cuda/nn_kernels.cu).

Every function launches exactly one kernel on the current torch CUDA stream, on tensors the
caller owns; nothing is allocated here.  Layouts follow Keras: NHWC activations, HWIO conv
kernels, (in, out) Dense kernels (README.md:292-298).
"""
from __future__ import annotations

import ctypes as C

import torch

from .._native import cuda_lib as nl

_declared = False
_DT = {torch.float32: 0, torch.bfloat16: 1}


class BnDesc(C.Structure):
    """Mirror of b2d_bn_desc (include/b200dist.h)."""
    _fields_ = [("dtype", C.c_int), ("M", C.c_int), ("C", C.c_int), ("relu", C.c_int), ("inference", C.c_int),
                ("eps", C.c_float), ("momentum", C.c_float), ("inv_lr", C.c_float),
                ("x", C.c_void_p), ("y", C.c_void_p), ("dy", C.c_void_p), ("dx", C.c_void_p),
                ("gamma", C.c_void_p), ("beta", C.c_void_p),
                ("moving_mean", C.c_void_p), ("moving_var", C.c_void_p),
                ("moving_mean_grad", C.c_void_p), ("moving_var_grad", C.c_void_p),
                ("stats", C.c_void_p), ("saved", C.c_void_p), ("red", C.c_void_p),
                ("dgamma", C.c_void_p), ("dbeta", C.c_void_p), ("residual", C.c_void_p)]


def _lib():
    global _declared
    lib = nl.lib()
    if not _declared:
        vp, i, f, sz, i64 = C.c_void_p, C.c_int, C.c_float, C.c_size_t, C.c_int64
        lib.b2d_conv_c1_pool_fwd.argtypes = [vp, vp, vp, vp, vp, i, i, i, i, i, vp]
        lib.b2d_conv_c1_pool_bwd.argtypes = [vp, vp, vp, vp, vp, i, i, i, i, i, vp]
        lib.b2d_dense_xent_head.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i, i, i, f, i, vp]
        lib.b2d_softmax_xent.argtypes = [vp, vp, vp, vp, i, i, f, vp]
        lib.b2d_maxpool2_fwd.argtypes = [vp, vp, vp, i, i, i, i, vp]
        lib.b2d_maxpool2_bwd.argtypes = [vp, vp, vp, vp, i, i, i, i, i, vp]
        lib.b2d_colsum.argtypes = [vp, vp, i, i, vp]
        lib.b2d_relu_bwd.argtypes = [vp, vp, vp, sz, vp]
        conv = [i] * 13
        lib.b2d_conv2d_fwd_direct.argtypes = [vp, vp, vp, vp, *conv, i, vp]
        lib.b2d_conv2d_dgrad_direct.argtypes = [vp, vp, vp, *conv, vp]
        lib.b2d_conv2d_wgrad_direct.argtypes = [vp, vp, vp, *conv, vp]
        lib.b2d_sgemm_simt.argtypes = [vp, i64, i64, vp, i64, i64, vp, i64, vp, i, i, i, i, i, vp]
        lib.b2d_sgd_apply.argtypes = [vp, vp, sz, sz, f, vp]
        lib.b2d_cast_f32_bf16.argtypes = [vp, vp, sz, vp]
        lib.b2d_bn_forward.argtypes = [C.POINTER(BnDesc), vp]
        lib.b2d_bn_backward.argtypes = [C.POINTER(BnDesc), vp]
        lib.b2d_add_act.argtypes = [i, vp, vp, vp, vp, sz, i, vp]
        lib.b2d_maxpool_fwd.argtypes = [i, vp, vp, vp, i, i, i, i, i, i, i, i, i, vp]
        lib.b2d_maxpool_bwd.argtypes = [i, vp, vp, vp, i, i, i, i, i, i, i, i, i, vp]
        lib.b2d_gap_fwd.argtypes = [i, vp, vp, i, i, i, vp]
        lib.b2d_gap_bwd.argtypes = [i, vp, vp, vp, i, i, i, vp]
        lib.b2d_im2col.argtypes = [i, i, vp, vp] + [i] * 13 + [vp]
        lib.b2d_cast.argtypes = [i, i, vp, vp, sz, vp]
        lib.b2d_colsum_t.argtypes = [i, vp, vp, i, i, vp]
        _declared = True
    return lib


def _s():
    return torch.cuda.current_stream().cuda_stream


def _p(t):
    return None if t is None else t.data_ptr()


def conv_c1_pool_fwd(x, w, b, y, code):
    """Conv2D(C_in=1, KxK valid) + bias + ReLU + MaxPool2x2.  x [N,H,W,1], w [K,K,1,Co]."""
    N, H, W, _ = x.shape
    KS, Co = w.shape[0], w.shape[3]
    nl.check(_lib().b2d_conv_c1_pool_fwd(_p(x), _p(w), _p(b), _p(y), _p(code), N, H, W, Co, KS, _s()),
             "conv_c1_pool_fwd")


def conv_c1_pool_bwd(x, dy, code, dw, db):
    """Accumulates dW [K,K,1,Co] and db [Co] from the pooled-output gradient dy."""
    N, H, W, _ = x.shape
    KS, Co = dw.shape[0], dw.shape[3]
    nl.check(_lib().b2d_conv_c1_pool_bwd(_p(x), _p(dy), _p(code), _p(dw), _p(db), N, H, W, Co, KS, _s()),
             "conv_c1_pool_bwd")


def dense_xent_head(h, w, b, labels, dh, dw, db, metrics, grad_scale, relu_mask, logits=None,
                    dh_colsum=None):
    """Dense(C<=32) + softmax cross-entropy + every gradient of both, one kernel.

    ``dh_colsum`` (optional, [K]) additionally accumulates the column sums of ``dh`` — the bias
    gradient of the layer that produced ``h``."""
    B, K = h.shape
    Cc = w.shape[1]
    nl.check(_lib().b2d_dense_xent_head(_p(h), _p(w), _p(b), _p(labels), _p(logits), _p(dh), _p(dw),
                                        _p(db), _p(dh_colsum), _p(metrics), B, K, Cc, float(grad_scale),
                                        int(relu_mask), _s()), "dense_xent_head")


def softmax_xent(logits, labels, dlogits, metrics, grad_scale):
    B, Cc = logits.shape
    nl.check(_lib().b2d_softmax_xent(_p(logits), _p(labels), _p(dlogits), _p(metrics), B, Cc,
                                     float(grad_scale), _s()), "softmax_xent")


def maxpool2_fwd(x, y, code):
    N, H, W, Cc = x.shape
    nl.check(_lib().b2d_maxpool2_fwd(_p(x), _p(y), _p(code), N, H, W, Cc, _s()), "maxpool2_fwd")


def maxpool2_bwd(dy, code, y_pooled, dx, act_relu):
    N, H, W, Cc = dx.shape
    nl.check(_lib().b2d_maxpool2_bwd(_p(dy), _p(code), _p(y_pooled), _p(dx), N, H, W, Cc,
                                     int(act_relu), _s()), "maxpool2_bwd")


def colsum(dy2d, db):
    M, N = dy2d.shape
    nl.check(_lib().b2d_colsum(_p(dy2d), _p(db), M, N, _s()), "colsum")


def relu_bwd(dy, y, dx):
    nl.check(_lib().b2d_relu_bwd(_p(dy), _p(y), _p(dx), dy.numel(), _s()), "relu_bwd")


def _conv_args(x_shape, w_shape, strides, padding, out_shape):
    """padding = (top, left); the output tensor's own height / width fix the bottom / right padding."""
    N, H, W, Ci = x_shape
    KH, KW, _, Co = w_shape
    return [N, H, W, Ci, Co, KH, KW, strides[0], strides[1], padding[0], padding[1], out_shape[1], out_shape[2]]


def conv2d_fwd_direct(x, w, b, y, strides=(1, 1), padding=(0, 0), relu=False):
    nl.check(_lib().b2d_conv2d_fwd_direct(_p(x), _p(w), _p(b), _p(y),
                                          *_conv_args(x.shape, w.shape, strides, padding, y.shape),
                                          int(relu), _s()), "conv2d_fwd_direct")


def conv2d_dgrad_direct(dy, w, dx, strides=(1, 1), padding=(0, 0)):
    nl.check(_lib().b2d_conv2d_dgrad_direct(_p(dy), _p(w), _p(dx),
                                            *_conv_args(dx.shape, w.shape, strides, padding, dy.shape), _s()),
             "conv2d_dgrad_direct")


def conv2d_wgrad_direct(x, dy, dw, strides=(1, 1), padding=(0, 0)):
    nl.check(_lib().b2d_conv2d_wgrad_direct(_p(x), _p(dy), _p(dw),
                                            *_conv_args(x.shape, dw.shape, strides, padding, dy.shape), _s()),
             "conv2d_wgrad_direct")


def sgemm_simt(a, b, d, *, trans_a=False, trans_b=False, bias=None, relu=False, accumulate=False):
    """D = op(A) @ op(B) on CUDA cores; operands are 2-D row-major fp32 tensors."""
    sam, sak = (a.stride(1), a.stride(0)) if trans_a else (a.stride(0), a.stride(1))
    sbk, sbn = (b.stride(1), b.stride(0)) if trans_b else (b.stride(0), b.stride(1))
    M, N = d.shape
    K = a.shape[0] if trans_a else a.shape[1]
    nl.check(_lib().b2d_sgemm_simt(_p(a), sam, sak, _p(b), sbk, sbn, _p(d), d.stride(0), _p(bias),
                                   M, N, K, int(relu), int(accumulate), _s()), "sgemm_simt")


def sgd_apply(w, g, n_params, lr):
    nl.check(_lib().b2d_sgd_apply(_p(w), _p(g), n_params, w.numel(), float(lr), _s()), "sgd_apply")


def cast_f32_bf16(src, dst):
    nl.check(_lib().b2d_cast_f32_bf16(_p(src), _p(dst), src.numel(), _s()), "cast_f32_bf16")


# ---- ResNet-family kernels (BASELINE.json config 4; csrc/cuda/resnet_kernels.cu) -------------
class BatchNorm:
    """One BatchNormalization layer instance over x2d [M, C] (fp32 or bf16 storage): owns the
    descriptor, so forward / backward are one C call each with no per-step host work.

    ``moving`` = (mean, var) fp32 [C] tensors; ``moving_grads`` = matching gradient-arena slots (the
    optimizer then performs the moving-average update, averaged over replicas) or None to update
    ``moving`` in place.  ``stats``/``saved``/``red`` are zero-initialised [2C] scratch tensors."""

    def __init__(self, x2d, y2d, gamma, beta, *, eps, momentum, relu, stats=None, saved=None, red=None,
                 moving=None, moving_grads=None, inv_lr=0.0, inference=False,
                 dy2d=None, dx2d=None, dgamma=None, dbeta=None, residual=None):
        M, Cc = x2d.shape
        d = BnDesc()
        d.dtype, d.M, d.C, d.relu, d.inference = _DT[x2d.dtype], M, Cc, int(relu), int(inference)
        d.eps, d.momentum, d.inv_lr = float(eps), float(momentum), float(inv_lr)
        d.x, d.y, d.dy, d.dx = _p(x2d), _p(y2d), _p(dy2d), _p(dx2d)
        d.gamma, d.beta = _p(gamma), _p(beta)
        if moving is not None:
            d.moving_mean, d.moving_var = _p(moving[0]), _p(moving[1])
        if moving_grads is not None:
            d.moving_mean_grad, d.moving_var_grad = _p(moving_grads[0]), _p(moving_grads[1])
        d.stats, d.saved, d.red, d.dgamma, d.dbeta = _p(stats), _p(saved), _p(red), _p(dgamma), _p(dbeta)
        d.residual = _p(residual)
        self.desc = d
        self._keep = [x2d, y2d, gamma, beta, stats, saved, red, moving, moving_grads, dy2d, dx2d, dgamma, dbeta, residual]
        self._lib = _lib()

    def forward(self):
        nl.check(self._lib.b2d_bn_forward(C.byref(self.desc), _s()), "bn_forward")

    def backward(self):
        nl.check(self._lib.b2d_bn_backward(C.byref(self.desc), _s()), "bn_backward")


def add_act(a, b, out, mode, mask=None):
    """mode 0: out = a + b; 1: relu(a + b); 2: (a + b) where mask > 0 else 0.  ``b`` may be None."""
    nl.check(_lib().b2d_add_act(_DT[a.dtype], _p(a), _p(b), _p(mask), _p(out), a.numel(), int(mode), _s()), "add_act")


def maxpool_fwd(x, y, code, k, s, p):
    N, H, W, Cc = x.shape
    nl.check(_lib().b2d_maxpool_fwd(_DT[x.dtype], _p(x), _p(y), _p(code), N, H, W, Cc, k, s, p, y.shape[1],
                                    y.shape[2], _s()), "maxpool_fwd")


def maxpool_bwd(dy, code, dx, k, s, p):
    N, H, W, Cc = dx.shape
    nl.check(_lib().b2d_maxpool_bwd(_DT[dx.dtype], _p(dy), _p(code), _p(dx), N, H, W, Cc, k, s, p, dy.shape[1],
                                    dy.shape[2], _s()), "maxpool_bwd")


def gap_fwd(x, y):
    """y [N, C] fp32 = mean over H, W of x [N, H, W, C] (fp32 or bf16)."""
    N, H, W, Cc = x.shape
    nl.check(_lib().b2d_gap_fwd(_DT[x.dtype], _p(x), _p(y), N, H * W, Cc, _s()), "gap_fwd")


def gap_bwd(dy, dx, mask=None):
    N, H, W, Cc = dx.shape
    nl.check(_lib().b2d_gap_bwd(_DT[dx.dtype], _p(dy), _p(mask), _p(dx), N, H * W, Cc, _s()), "gap_bwd")


def im2col(x, col, kernel, strides, padding, out_hw):
    """col[pixel, :] = flattened (kh, kw, ci) patch of x for every output pixel; col is [N*OH*OW, kpad]."""
    N, H, W, Ci = x.shape
    nl.check(_lib().b2d_im2col(_DT[x.dtype], _DT[col.dtype], _p(x), _p(col), N, H, W, Ci, kernel[0], kernel[1],
                               strides[0], strides[1], padding[0], padding[1], out_hw[0], out_hw[1], col.shape[1],
                               _s()), "im2col")


def cast(src, dst):
    """dst = src converted between fp32 and bf16 (same element count)."""
    nl.check(_lib().b2d_cast(_DT[src.dtype], _DT[dst.dtype], _p(src), _p(dst), src.numel(), _s()), "cast")


def colsum_t(dy2d, db):
    """db += column sums of dy2d [M, C] (fp32 or bf16 storage, C a multiple of 4 / 8)."""
    M, Cc = dy2d.shape
    nl.check(_lib().b2d_colsum_t(_DT[dy2d.dtype], _p(dy2d), _p(db), M, Cc, _s()), "colsum_t")
