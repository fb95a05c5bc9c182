"""CPU backend: lowers ops to the C++ kernels in libb200dist_cpu.so (oracle + README config 1)."""
from __future__ import annotations

import torch

from ..ops import cpu


class CpuBackend:
    name = "cpu"
    device = torch.device("cpu")
    supports_conv_c1_pool = False

    def alloc(self, shape, dtype=torch.float32):
        return torch.zeros(shape, dtype=dtype)

    def supports_fused_head(self, k_in, units):
        return False

    def describe_conv(self, layer):
        return "C++ direct conv"

    def describe_dense(self, b, k, n):
        return "C++ gemm"

    # every method returns a zero-argument callable that runs the op on fixed buffers
    def conv2d_fwd(self, x, w, b, y, strides, pad, relu):
        return lambda: cpu.conv2d_fwd(x, w, b, y, strides, pad, relu)

    def conv2d_wgrad(self, x, dy, dw, db, strides, pad):
        return lambda: cpu.conv2d_wgrad(x, dy, dw, db, strides, pad)

    def conv2d_dgrad(self, dy, w, dx, strides, pad, mask=None):
        def run():
            cpu.conv2d_dgrad(dy, w, dx, strides, pad)
            if mask is not None:
                cpu.relu_bwd(dx, mask, dx)
        return run

    def maxpool2_fwd(self, x, y, code):
        return lambda: cpu.maxpool2_fwd(x, y, code)

    def maxpool2_bwd(self, dy, code, y, dx, act_relu):
        return lambda: cpu.maxpool2_bwd(dy, code, y, dx, act_relu)

    def dense_fwd(self, x, w, b, y, relu):
        return lambda: cpu.gemm(x, w, y, bias=b, relu=relu)

    def dense_wgrad(self, x, dy, dw, db):
        def run():
            cpu.gemm(x, dy, dw, trans_a=True, accumulate=True)
            if db is not None:
                cpu.colsum(dy, db)
        return run

    def dense_dgrad(self, dy, w, dx, mask=None):
        def run():
            cpu.gemm(dy, w, dx, trans_b=True)
            if mask is not None:
                cpu.relu_bwd(dx, mask, dx)
        return run

    def softmax_xent(self, logits, labels, dlogits, metrics, scale):
        return lambda: cpu.softmax_xent(logits, labels, dlogits, metrics, scale)

    def relu_bwd(self, dy, y, dx):
        return lambda: cpu.relu_bwd(dy, y, dx)

    def __getattr__(self, name):
        if name in ("bn_fwd", "bn_bwd", "add_act", "copy", "maxpool_fwd", "maxpool_bwd", "gap_fwd", "gap_bwd"):
            raise NotImplementedError(f"{name}: BatchNormalization / Residual / general pooling layers run on the "
                                      "CUDA backend only (the CPU backend covers the reference's layer set)")
        raise AttributeError(name)

    def sgd_apply(self, params, grads, n_params, lr):
        return lambda: cpu.sgd_apply(params, grads, n_params, lr)
