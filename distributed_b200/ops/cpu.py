"""C++ CPU ops on torch CPU tensors (csrc/cpu/ops.cpp): the oracle and the CPU backend.

Same signatures and layouts as :mod:`distributed_b200.ops.nn` (NHWC / HWIO / (in, out)).
"""
from __future__ import annotations

from .._native import cpu_lib


def _p(t):
    return None if t is None else t.data_ptr()


def _conv_args(x_shape, w_shape, strides, padding, out_shape):
    """padding = (top, left); the output tensor's own height / width fix the bottom / right padding."""
    N, H, W, Ci = x_shape
    KH, KW, _, Co = w_shape
    return [N, H, W, Ci, Co, KH, KW, strides[0], strides[1], padding[0], padding[1], out_shape[1], out_shape[2]]


def conv2d_fwd(x, w, b, y, strides=(1, 1), padding=(0, 0), relu=False):
    cpu_lib.lib().b2c_conv2d_fwd(_p(x), _p(w), _p(b), _p(y), *_conv_args(x.shape, w.shape, strides, padding, y.shape),
                                 int(relu))


def conv2d_wgrad(x, dy, dw, db, strides=(1, 1), padding=(0, 0)):
    cpu_lib.lib().b2c_conv2d_wgrad(_p(x), _p(dy), _p(dw), _p(db),
                                   *_conv_args(x.shape, dw.shape, strides, padding, dy.shape))


def conv2d_dgrad(dy, w, dx, strides=(1, 1), padding=(0, 0)):
    cpu_lib.lib().b2c_conv2d_dgrad(_p(dy), _p(w), _p(dx), *_conv_args(dx.shape, w.shape, strides, padding, dy.shape))


def maxpool2_fwd(x, y, code):
    N, H, W, C = x.shape
    cpu_lib.lib().b2c_maxpool2_fwd(_p(x), _p(y), _p(code), N, H, W, C)


def maxpool2_bwd(dy, code, y_pooled, dx, act_relu):
    N, H, W, C = dx.shape
    cpu_lib.lib().b2c_maxpool2_bwd(_p(dy), _p(code), _p(y_pooled), _p(dx), N, H, W, C, int(act_relu))


def gemm(a, b, d, *, trans_a=False, trans_b=False, bias=None, relu=False, accumulate=False):
    sam, sak = (a.stride(1), a.stride(0)) if trans_a else (a.stride(0), a.stride(1))
    sbk, sbn = (b.stride(1), b.stride(0)) if trans_b else (b.stride(0), b.stride(1))
    M, N = d.shape
    K = a.shape[0] if trans_a else a.shape[1]
    cpu_lib.lib().b2c_gemm(_p(a), sam, sak, _p(b), sbk, sbn, _p(d), d.stride(0), _p(bias), M, N, K,
                           int(relu), int(accumulate))


def softmax_xent(logits, labels, dlogits, metrics, grad_scale):
    B, C = logits.shape
    cpu_lib.lib().b2c_softmax_xent(_p(logits), _p(labels), _p(dlogits), _p(metrics), B, C, float(grad_scale))


def relu_bwd(dy, y, dx):
    cpu_lib.lib().b2c_relu_bwd(_p(dy), _p(y), _p(dx), dy.numel())


def colsum(dy2d, db):
    M, N = dy2d.shape
    cpu_lib.lib().b2c_colsum(_p(dy2d), _p(db), M, N)


def sgd_apply(w, g, n_params, lr):
    cpu_lib.lib().b2c_sgd_apply(_p(w), _p(g), n_params, w.numel(), float(lr))


# ---- ResNet-family layers (csrc/cpu/resnet_ops.cpp) -------------------------------------------
def bn_fwd(x2d, gamma, beta, y2d, saved, moving, moving_grads, eps, momentum, inv_lr, relu, inference,
           residual=None):
    M, C = x2d.shape
    mm, mv = moving if moving is not None else (None, None)
    gm, gv = moving_grads if moving_grads is not None else (None, None)
    cpu_lib.lib().b2c_bn_fwd(_p(x2d), _p(gamma), _p(beta), _p(y2d), _p(saved), _p(mm), _p(mv), _p(gm), _p(gv),
                             M, C, float(eps), float(momentum), float(inv_lr), int(relu), int(inference), _p(residual))


def bn_bwd(x2d, y2d, dy2d, gamma, saved, dx2d, dgamma, dbeta, relu):
    M, C = x2d.shape
    cpu_lib.lib().b2c_bn_bwd(_p(x2d), _p(y2d), _p(dy2d), _p(gamma), _p(saved), _p(dx2d), _p(dgamma), _p(dbeta),
                             M, C, int(relu))


def add_act(a, b, out, mode, mask=None):
    cpu_lib.lib().b2c_add_act(_p(a), _p(b), _p(mask), _p(out), a.numel(), int(mode))


def maxpool_fwd(x, y, code, k, s, p):
    N, H, W, C = x.shape
    cpu_lib.lib().b2c_maxpool_fwd(_p(x), _p(y), _p(code), N, H, W, C, k, s, p, y.shape[1], y.shape[2])


def maxpool_bwd(dy, code, dx, k, s, p):
    N, H, W, C = dx.shape
    cpu_lib.lib().b2c_maxpool_bwd(_p(dy), _p(code), _p(dx), N, H, W, C, k, s, p, dy.shape[1], dy.shape[2])


def gap_fwd(x, y):
    N, H, W, C = x.shape
    cpu_lib.lib().b2c_gap_fwd(_p(x), _p(y), N, H * W, C)


def gap_bwd(dy, dx, mask=None):
    N, H, W, C = dx.shape
    cpu_lib.lib().b2c_gap_bwd(_p(dy), _p(mask), _p(dx), N, H * W, C)
