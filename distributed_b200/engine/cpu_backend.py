"""CPU backend: lowers ops to the C++ kernels in libb200dist_cpu.so (oracle + README config 1)."""
from __future__ import annotations

import torch

from ..ops import cpu


class CpuBackend:
    name = "cpu"
    device = torch.device("cpu")
    supports_conv_c1_pool = False
    supports_bn_residual = True
    act_dtype = torch.float32

    def alloc(self, shape, dtype=torch.float32):
        return torch.zeros(shape, dtype=dtype or torch.float32)

    def supports_fused_head(self, k_in, units):
        return False

    def describe_conv(self, layer):
        k = layer.kernel_size[0] * layer.kernel_size[1] * layer.input_shape[2]
        return "C++ im2col + tiled GEMM" if k >= 32 else "C++ direct conv"

    def describe_dense(self, b, k, n):
        return "C++ gemm"

    # every method returns a zero-argument callable that runs the op on fixed buffers
    def conv2d_fwd(self, x, w, b, y, strides, pad, relu):
        return lambda: cpu.conv2d_fwd(x, w, b, y, strides, pad, relu)

    def conv2d_wgrad(self, x, dy, dw, db, strides, pad):
        return lambda: cpu.conv2d_wgrad(x, dy, dw, db, strides, pad)

    def dgrad_addend_ok(self, dx_shape, w_shape, strides, pad, out_hw, dtype):
        return tuple(strides) == (1, 1)         # mirrors the CUDA backend's fused residual gradient sum

    def conv2d_dgrad(self, dy, w, dx, strides, pad, mask=None, addend=None):
        def run():
            cpu.conv2d_dgrad(dy, w, dx, strides, pad)
            if addend is not None:
                cpu.add_act(dx, addend, dx, 2 if mask is not None else 0, mask)
            elif mask is not None:
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

    # ---- ResNet-family layers (fp32; the dtype policy only affects the CUDA backend) --------------
    def batchnorm(self, x2d, y2d, gamma, beta, *, eps, momentum, relu, stats=None, saved=None, red=None,
                  moving=None, moving_grads=None, inv_lr=0.0, inference=False, dy2d=None, dx2d=None,
                  dgamma=None, dbeta=None, residual=None):
        def fwd():
            cpu.bn_fwd(x2d, gamma, beta, y2d, saved, moving, moving_grads, eps, momentum, inv_lr, relu, inference,
                       residual)

        def bwd():      # with a residual input the block's output gradient arrives already masked
            cpu.bn_bwd(x2d, y2d, dy2d, gamma, saved, dx2d, dgamma, dbeta, relu and residual is None)
        return fwd, bwd

    def add_act(self, a, b, out, mode, mask=None):
        return lambda: cpu.add_act(a, b, out, mode, mask)

    def cast(self, src, dst):
        return lambda: dst.copy_(src)

    def maxpool_fwd(self, x, y, code, k, s, p):
        return lambda: cpu.maxpool_fwd(x, y, code, k, s, p)

    def maxpool_bwd(self, dy, code, dx, k, s, p):
        return lambda: cpu.maxpool_bwd(dy, code, dx, k, s, p)

    def gap_fwd(self, x, y):
        return lambda: cpu.gap_fwd(x, y)

    def gap_bwd(self, dy, dx, mask=None):
        return lambda: cpu.gap_bwd(dy, dx, mask)

    def sgd_apply(self, params, grads, n_params, lr):
        return lambda: cpu.sgd_apply(params, grads, n_params, lr)
