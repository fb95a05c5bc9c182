"""CUDA backend: lowers ops to the sm_100a kernels in libb200dist_cuda.so.

Dense layers go to the tcgen05 GEMM core whenever their shapes meet TMA's 16-byte row
alignment; the tiny/unaligned remainder (e.g. an unfused Dense(10)) uses the CUDA-core GEMM.
Every callable launches on the *current* torch stream, so the executor can capture them.
"""
from __future__ import annotations

import os

import torch

from ..ops import nn
from ..ops.gemm import GemmPlan

_HEAD_MAX_C = 32
_HEAD_SMEM = 200 * 1024


class CudaBackend:
    name = "cuda"
    supports_conv_c1_pool = True

    def __init__(self, device):
        self.device = torch.device(device)
        self.plans = []     # keep GemmPlans (TMA descriptors) alive

    def alloc(self, shape, dtype=torch.float32):
        return torch.zeros(shape, dtype=dtype, device=self.device)

    def supports_fused_head(self, k_in, units):
        smem = (2 * k_in * units + 17 * k_in + 8 * _HEAD_MAX_C + 16) * 4
        return units <= _HEAD_MAX_C and smem <= _HEAD_SMEM

    @staticmethod
    def _tc_ok(*dims):
        # B200DIST_DENSE=simt forces the exact-fp32 CUDA-core GEMM (debugging / bit-level checks)
        if os.environ.get("B200DIST_DENSE", "") == "simt":
            return False
        return all(d % 4 == 0 for d in dims)

    def describe_conv(self, layer):
        from ..ops import conv as tc
        ci = layer.input_shape[2]
        x_shape = (1, *layer.input_shape)
        w_shape = (*layer.kernel_size, ci, layer.filters)
        if self._conv_tc_ok(x_shape, w_shape, layer.strides, layer.pad_amounts(), out_hw=layer.output_shape[:2]):
            return "tcgen05 tf32 implicit GEMM (4-D TMA taps)"
        return "direct conv (CUDA cores)"

    @staticmethod
    def _conv_tc_ok(x_shape, w_shape, strides, pad, dgrad=False, out_hw=None):
        from ..ops import conv as tc
        if os.environ.get("B200DIST_CONV", "") == "direct":
            return False
        if dgrad and out_hw is not None and tuple(strides) == (1, 1):
            # the unit-stride data gradient assumes symmetric padding
            if (x_shape[1] + 2 * pad[0] - w_shape[0] + 1, x_shape[2] + 2 * pad[1] - w_shape[1] + 1) != tuple(out_hw):
                return False
        return tc.tc_supported(tuple(x_shape), tuple(w_shape), pad, torch.float32, dgrad=dgrad, strides=strides,
                               out_hw=out_hw)

    def describe_dense(self, b, k, n):
        return "tcgen05 tf32 GEMM" if self._tc_ok(k, n) else "CUDA-core GEMM"

    def _plan(self, *a, **kw):
        p = GemmPlan(*a, **kw)
        self.plans.append(p)
        return p.run

    # ---- fused first layer -----------------------------------------------------------
    def conv_c1_pool_fwd(self, x, w, b, y, code):
        return lambda: nn.conv_c1_pool_fwd(x, w, b, y, code)

    def conv_c1_pool_bwd(self, x, dy, code, dw, db):
        return lambda: nn.conv_c1_pool_bwd(x, dy, code, dw, db)

    # ---- general conv (direct fallback; implicit-GEMM path plugs in here) --------------
    def conv2d_fwd(self, x, w, b, y, strides, pad, relu):
        from ..ops import conv as tc
        if self._conv_tc_ok(x.shape, w.shape, strides, pad, out_hw=y.shape[1:3]):
            plan = tc.ConvFwdPlan(x, w, y, pad, bias=b, relu=relu, stride=strides[0])
            self.plans.append(plan)
            return plan.run
        if self._im2col_ok(x, w, y):
            # small C_in (e.g. the 3-channel ResNet stem): patch matrix once, then tensor-core GEMMs
            col = self._col_for(x, w, y, strides, pad)
            k = w.shape[0] * w.shape[1] * w.shape[2]
            gemm = self._plan(col[:, :k], w.view(k, w.shape[3]), y.view(-1, y.shape[3]), b_mn_major=True, bias=b,
                              relu=relu)
            kh_kw, st, pd, ohw = (w.shape[0], w.shape[1]), tuple(strides), tuple(pad), (y.shape[1], y.shape[2])

            def run():
                nn.im2col(x, col, kh_kw, st, pd, ohw)
                gemm()
            return run
        return lambda: nn.conv2d_fwd_direct(x, w, b, y, strides, pad, relu)

    def _im2col_ok(self, x, w, y):
        k = w.shape[0] * w.shape[1] * w.shape[2]
        return (os.environ.get("B200DIST_CONV", "") != "direct" and (x.shape[3] * 4) % 128 != 0 and k >= 32
                and w.shape[3] % 4 == 0 and y.numel() // y.shape[3] >= 1024)

    def _col_for(self, x, w, y, strides, pad):
        """One patch matrix per (input tensor, geometry): written by the forward, reused by wgrad."""
        k = w.shape[0] * w.shape[1] * w.shape[2]
        key = (x.data_ptr(), tuple(w.shape[:3]), tuple(strides), tuple(pad), y.shape[1], y.shape[2])
        cache = self.__dict__.setdefault("_cols", {})
        if key not in cache:
            cache[key] = self.alloc((y.numel() // y.shape[3], (k + 3) // 4 * 4))
        return cache[key]

    def conv2d_wgrad(self, x, dy, dw, db, strides, pad):
        from ..ops import conv as tc
        dy2 = dy.view(-1, dy.shape[-1])
        if (self._conv_tc_ok(x.shape, dw.shape, strides, pad, out_hw=dy.shape[1:3]) and x.shape[3] % 4 == 0
                and dw.shape[3] % 4 == 0):
            plan = tc.ConvWgradPlan(x, dy, dw, pad, stride=strides[0])
            self.plans.append(plan)
            wgrad = plan.run
        elif self._im2col_ok(x, dw, dy):
            col = self._col_for(x, dw, dy, strides, pad)        # filled by this step's forward pass
            k = dw.shape[0] * dw.shape[1] * dw.shape[2]
            # the padding columns of col are excluded by the view; its row pitch stays 16-byte aligned
            wgrad = self._plan(col[:, :k], dy2, dw.view(k, dw.shape[3]), a_mn_major=True, b_mn_major=True,
                               accumulate=True)
        else:
            def wgrad():
                nn.conv2d_wgrad_direct(x, dy, dw, strides, pad)

        def run():
            wgrad()
            if db is not None:
                nn.colsum(dy2, db)
        return run

    def conv2d_dgrad(self, dy, w, dx, strides, pad, mask=None):
        from ..ops import conv as tc
        if (strides[0] == strides[1] and strides[0] > 1 and os.environ.get("B200DIST_CONV", "") != "direct"
                and tc.dgrad_strided_supported(dx.shape, w.shape, strides[0], torch.float32)):
            plan = tc.ConvDgradStridedPlan(dy, w, dx, pad, strides[0], mask=mask)
            self.plans.append(plan)
            return plan.run
        if self._conv_tc_ok(dx.shape, w.shape, strides, pad, dgrad=True, out_hw=dy.shape[1:3]):
            plan = tc.ConvDgradPlan(dy, w, dx, pad, mask=mask)     # ReLU mask fused in the epilogue
            self.plans.append(plan)
            return plan.run

        def run():
            nn.conv2d_dgrad_direct(dy, w, dx, strides, pad)
            if mask is not None:
                nn.relu_bwd(dx, mask, dx)
        return run

    def maxpool2_fwd(self, x, y, code):
        return lambda: nn.maxpool2_fwd(x, y, code)

    def maxpool2_bwd(self, dy, code, y, dx, act_relu):
        return lambda: nn.maxpool2_bwd(dy, code, y, dx, act_relu)

    # ---- dense ------------------------------------------------------------------------
    def dense_fwd(self, x, w, b, y, relu):
        k, n = w.shape
        if self._tc_ok(k, n):
            return self._plan(x, w, y, b_mn_major=True, bias=b, relu=relu)
        return lambda: nn.sgemm_simt(x, w, y, bias=b, relu=relu)

    def dense_wgrad(self, x, dy, dw, db):
        k, n = dw.shape
        if self._tc_ok(k, n):
            gemm = self._plan(x, dy, dw, a_mn_major=True, b_mn_major=True, accumulate=True)
        else:
            def gemm():
                nn.sgemm_simt(x, dy, dw, trans_a=True, accumulate=True)

        def run():
            gemm()
            if db is not None:
                nn.colsum(dy, db)
        return run

    def dense_dgrad(self, dy, w, dx, mask=None):
        k, n = w.shape
        if self._tc_ok(k, n):
            return self._plan(dy, w, dx, mask=mask)        # dx = dy . W^T, W rows are K-major
        def run():
            nn.sgemm_simt(dy, w, dx, trans_b=True)
            if mask is not None:
                nn.relu_bwd(dx, mask, dx)
        return run

    def dense_xent_head(self, h, w, b, labels, dh, dw, db, metrics, scale, relu_mask, logits, dh_colsum=None):
        return lambda: nn.dense_xent_head(h, w, b, labels, dh, dw, db, metrics, scale, relu_mask, logits,
                                          dh_colsum)

    # ---- ResNet-family ops --------------------------------------------------------------
    def bn_fwd(self, x2d, gamma, beta, stats, saved, moving, y2d, eps, momentum, relu):
        return lambda: nn.bn_fwd(x2d, gamma, beta, stats, saved, moving, y2d, eps, momentum, relu)

    def bn_bwd(self, x2d, y2d, dy2d, gamma, saved, red, dx2d, dgamma, dbeta, relu):
        return lambda: nn.bn_bwd(x2d, y2d, dy2d, gamma, saved, red, dx2d, dgamma, dbeta, relu)

    def add_act(self, a, b, out, relu):
        return lambda: nn.add_act(a, b, out, relu)

    def relu_bwd(self, dy, y, dx):
        return lambda: nn.relu_bwd(dy, y, dx)

    def copy(self, src, dst):
        return lambda: dst.copy_(src)          # contiguous D2D: a memcpy node, not a kernel

    def maxpool_fwd(self, x, y, code, k, s, p):
        return lambda: nn.maxpool_fwd(x, y, code, k, s, p)

    def maxpool_bwd(self, dy, code, dx, k, s, p):
        return lambda: nn.maxpool_bwd(dy, code, dx, k, s, p)

    def gap_fwd(self, x, y):
        return lambda: nn.gap_fwd(x, y)

    def gap_bwd(self, dy, dx):
        return lambda: nn.gap_bwd(dy, dx)

    def softmax_xent(self, logits, labels, dlogits, metrics, scale):
        return lambda: nn.softmax_xent(logits, labels, dlogits, metrics, scale)

    def sgd_apply(self, params, grads, n_params, lr):
        return lambda: nn.sgd_apply(params, grads, n_params, lr)
