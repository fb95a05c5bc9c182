"""CUDA backend: lowers ops to the sm_100a kernels in libb200dist_cuda.so.

Dense layers go to the tcgen05 GEMM core whenever their shapes meet TMA's 16-byte row
alignment; the tiny/unaligned remainder (e.g. an unfused Dense(10)) uses the CUDA-core GEMM.
Every callable launches on the *current* torch stream, so the executor can capture them.

``act_dtype`` selects the storage type of the convolutional part of the network: fp32 (tf32
tensor-core math) or bf16 (the ``mixed_bfloat16`` policy: bf16 activations, activation gradients
and a bf16 shadow of the weights, fp32 master weights / weight gradients / statistics).
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
    supports_bn_residual = True

    def __init__(self, device, act_dtype=torch.float32):
        self.device = torch.device(device)
        self.act_dtype = act_dtype      # storage type of 4-D activations (fp32, or bf16 = mixed precision)
        self.plans = []                 # keep GemmPlans (TMA descriptors) alive
        self.scattered = []             # arena ranges reduce-scattered by their producing GEMM (see dense_wgrad)
        self._params = self._shadow = None
        self._cols, self._casts = {}, {}

    def alloc(self, shape, dtype=torch.float32):
        return torch.zeros(shape, dtype=dtype or torch.float32, device=self.device)

    # ---- mixed precision: bf16 shadow of the fp32 master weights ---------------------------
    def begin_step(self, params):
        """First launch of a step under the bf16 policy: refresh the bf16 copy of the whole weight
        arena (one streaming kernel; the fp32 masters are what the optimizer updates)."""
        if self.act_dtype != torch.bfloat16:
            return None
        if self._params is None or self._params.data_ptr() != params.data_ptr():
            self._params = params
            self._shadow = torch.zeros(params.numel(), dtype=torch.bfloat16, device=self.device)
        shadow = self._shadow
        return lambda: nn.cast(params, shadow)

    def _w_for(self, w, dtype):
        """The weight tensor in the operand type of a `dtype` convolution."""
        if w.dtype == dtype:
            return w
        if self._params is None:
            raise RuntimeError("bf16 convolution requested before begin_step() bound the weight arena")
        off = (w.data_ptr() - self._params.data_ptr()) // 4
        if not 0 <= off <= self._params.numel() - w.numel():
            raise RuntimeError("bf16 convolution weight does not live in the bound weight arena")
        return self._shadow[off:off + w.numel()].view(w.shape)

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
        ci = layer.input_shape[2]
        x_shape = (1, *layer.input_shape)
        w_shape = (*layer.kernel_size, ci, layer.filters)
        kind = "bf16" if self.act_dtype == torch.bfloat16 else "tf32"
        if self._conv_tc_ok(x_shape, w_shape, layer.strides, layer.pad_amounts(), out_hw=layer.output_shape[:2],
                            dtype=self.act_dtype):
            return f"tcgen05 {kind} implicit GEMM (4-D TMA taps)"
        direct = os.environ.get("B200DIST_CONV", "") == "direct"
        if not direct and (self.act_dtype == torch.bfloat16 or (ci * 4) % 128 != 0):
            return f"im2col + tcgen05 {kind} GEMM"
        return "direct conv (CUDA cores)"

    @staticmethod
    def _conv_tc_ok(x_shape, w_shape, strides, pad, dgrad=False, out_hw=None, dtype=torch.float32):
        from ..ops import conv as tc
        if os.environ.get("B200DIST_CONV", "") == "direct":
            return False
        if dgrad and out_hw is not None and tuple(strides) == (1, 1):
            # the unit-stride data gradient assumes symmetric padding
            if (x_shape[1] + 2 * pad[0] - w_shape[0] + 1, x_shape[2] + 2 * pad[1] - w_shape[1] + 1) != tuple(out_hw):
                return False
        return tc.tc_supported(tuple(x_shape), tuple(w_shape), pad, dtype, dgrad=dgrad, strides=strides,
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

    # ---- general conv: implicit GEMM on tcgen05, im2col + GEMM for small C_in, direct fallback ----
    @staticmethod
    def _unsupported_bf16(what, x, w):
        raise NotImplementedError(
            f"mixed_bfloat16: {what} of a convolution with input {tuple(x.shape)} and kernel {tuple(w.shape)} has no "
            "bf16 kernel — interior Conv2D layers need C_in % 64 == 0 and C_out % 8 == 0 (the first layer may have "
            "any C_in); use the float32 policy for this model")

    def conv2d_fwd(self, x, w, b, y, strides, pad, relu):
        from ..ops import conv as tc
        dt = y.dtype
        wk = self._w_for(w, dt)
        if x.dtype == dt and self._conv_tc_ok(x.shape, w.shape, strides, pad, out_hw=y.shape[1:3], dtype=dt):
            plan = tc.ConvFwdPlan(x, wk, y, pad, bias=b, relu=relu, stride=strides[0])
            self.plans.append(plan)
            return plan.run
        if self._im2col_ok(x, w, y):
            # small C_in (e.g. the 3-channel ResNet stem): patch matrix once, then tensor-core GEMMs
            col = self._col_for(x, w, y, strides, pad)
            k = w.shape[0] * w.shape[1] * w.shape[2]
            gemm = self._plan(col[:, :k], wk.view(k, w.shape[3]), y.view(-1, y.shape[3]), b_mn_major=True, bias=b,
                              relu=relu)
            kh_kw, st, pd, ohw = (w.shape[0], w.shape[1]), tuple(strides), tuple(pad), (y.shape[1], y.shape[2])

            def run():
                nn.im2col(x, col, kh_kw, st, pd, ohw)
                gemm()
            return run
        if dt != torch.float32:
            self._unsupported_bf16("the forward pass", x, w)
        return lambda: nn.conv2d_fwd_direct(x, w, b, y, strides, pad, relu)

    def _im2col_ok(self, x, w, y):
        k = w.shape[0] * w.shape[1] * w.shape[2]
        if os.environ.get("B200DIST_CONV", "") == "direct" or k < 32 or y.numel() // y.shape[3] < 1024:
            return False
        if y.dtype == torch.bfloat16:       # every convolution the implicit-GEMM path does not take
            return w.shape[3] % 8 == 0
        return (x.shape[3] * 4) % 128 != 0 and w.shape[3] % 4 == 0

    def _col_for(self, x, w, y, strides, pad):
        """One patch matrix per (input tensor, geometry): written by the forward, reused by wgrad."""
        k = w.shape[0] * w.shape[1] * w.shape[2]
        key = (x.data_ptr(), tuple(w.shape[:3]), tuple(strides), tuple(pad), y.shape[1], y.shape[2], y.dtype)
        if key not in self._cols:
            vec = 4 if y.dtype == torch.float32 else 8
            self._cols[key] = self.alloc((y.numel() // y.shape[3], (k + vec - 1) // vec * vec), y.dtype)
        return self._cols[key]

    def conv2d_wgrad(self, x, dy, dw, db, strides, pad):
        from ..ops import conv as tc
        dt = dy.dtype
        vec = 4 if dt == torch.float32 else 8
        dy2 = dy.view(-1, dy.shape[-1])
        pointwise = (tuple(dw.shape[:2]) == (1, 1) and tuple(strides) == (1, 1) and tuple(pad) == (0, 0)
                     and x.dtype == dt and x.shape[3] % vec == 0 and dw.shape[3] % vec == 0
                     and os.environ.get("B200DIST_CONV", "") != "direct"
                     and os.environ.get("B200DIST_WGRAD_1X1", "gemm") == "gemm")
        if pointwise:
            # a 1x1 / stride-1 filter gradient is a plain GEMM over the pixels: dW[ci, co] += x^T . dy,
            # split along the pixels, partial tiles added with 16-byte vector reductions
            wgrad = self._plan(x.view(-1, x.shape[3]), dy2, dw.view(dw.shape[2], dw.shape[3]), a_mn_major=True,
                               b_mn_major=True, accumulate=True)
        elif (x.dtype == dt and self._conv_tc_ok(x.shape, dw.shape, strides, pad, out_hw=dy.shape[1:3], dtype=dt)
                and x.shape[3] % vec == 0 and dw.shape[3] % vec == 0):
            plan = tc.ConvWgradPlan(x, dy, dw, pad, stride=strides[0])
            self.plans.append(plan)
            wgrad = plan.run
        elif self._im2col_ok(x, dw, dy):
            col = self._col_for(x, dw, dy, strides, pad)        # filled by this step's forward pass
            k = dw.shape[0] * dw.shape[1] * dw.shape[2]
            # the padding columns of col are excluded by the view; its row pitch stays 16-byte aligned
            wgrad = self._plan(col[:, :k], dy2, dw.view(k, dw.shape[3]), a_mn_major=True, b_mn_major=True,
                               accumulate=True)
        elif dt != torch.float32:
            self._unsupported_bf16("the filter gradient", x, dw)
        else:
            def wgrad():
                nn.conv2d_wgrad_direct(x, dy, dw, strides, pad)

        def run():
            wgrad()
            if db is not None:
                if dt == torch.float32:
                    nn.colsum(dy2, db)
                else:
                    nn.colsum_t(dy2, db)
        return run

    def dgrad_addend_ok(self, dx_shape, w_shape, strides, pad, out_hw, dtype):
        """True when the data gradient of this convolution can add another tensor in its epilogue."""
        return (tuple(strides) == (1, 1) and os.environ.get("B200DIST_FUSE_SUM", "1") != "0"
                and self._conv_tc_ok(dx_shape, w_shape, strides, pad, dgrad=True, out_hw=out_hw, dtype=dtype))

    def conv2d_dgrad(self, dy, w, dx, strides, pad, mask=None, addend=None):
        from ..ops import conv as tc
        dt = dx.dtype
        wk = self._w_for(w, dt)
        if addend is not None:
            assert self.dgrad_addend_ok(dx.shape, w.shape, strides, pad, dy.shape[1:3], dt)
            plan = tc.ConvDgradPlan(dy, wk, dx, pad, mask=mask, addend=addend)
            self.plans.append(plan)
            return plan.run
        if (strides[0] == strides[1] and strides[0] > 1 and os.environ.get("B200DIST_CONV", "") != "direct"
                and tc.dgrad_strided_supported(dx.shape, w.shape, strides[0], dt)):
            plan = tc.ConvDgradStridedPlan(dy, wk, dx, pad, strides[0], mask=mask)
            self.plans.append(plan)
            return plan.run
        if self._conv_tc_ok(dx.shape, w.shape, strides, pad, dgrad=True, out_hw=dy.shape[1:3], dtype=dt):
            plan = tc.ConvDgradPlan(dy, wk, dx, pad, mask=mask)     # ReLU mask fused in the epilogue
            self.plans.append(plan)
            return plan.run
        if dt != torch.float32:
            self._unsupported_bf16("the data gradient", dx, w)

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

    # GEMM -> reduce-scatter (set by the trainer before lowering a step whose update is overlapped):
    # ``scatter_target`` = {"bases": gradient arena of every replica, "local": this replica's arena tensor,
    # "rank", "world", "min_floats"}; ``scattered`` collects the arena ranges whose weight-gradient GEMM adds
    # its result straight into the owning replica's arena.
    scatter_target = None

    def dense_wgrad(self, x, dy, dw, db):
        k, n = dw.shape
        if self._tc_ok(k, n):
            plan = GemmPlan(x, dy, dw, a_mn_major=True, b_mn_major=True, accumulate=True)
            self.plans.append(plan)
            gemm = plan.run
            st = self.scatter_target
            if (st is not None and plan.geometry.splits == 1 and not plan.geometry.pair
                    and dw.numel() >= st["min_floats"]
                    and dw.numel() % 4 == 0 and dw.is_contiguous()):
                off = (dw.data_ptr() - st["local"].data_ptr()) // 4
                world = st["world"]
                per = -(-(dw.numel() // 4) // world) * 4           # the all-reduce kernel's shard size, in floats
                if 0 <= off and off + dw.numel() <= st["local"].numel() and off % 4 == 0:
                    plan.set_scatter(st["bases"], off, off, per, st["rank"])
                    self.scattered.append((off, off + dw.numel()))
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

    # ---- ResNet-family ops (fp32 or bf16 storage) -------------------------------------------
    def batchnorm(self, x2d, y2d, gamma, beta, **kw):
        bn = nn.BatchNorm(x2d, y2d, gamma, beta, **kw)
        self.plans.append(bn)
        return bn.forward, bn.backward

    def add_act(self, a, b, out, mode, mask=None):
        return lambda: nn.add_act(a, b, out, mode, mask)

    def relu_bwd(self, dy, y, dx):
        if dy.dtype == torch.float32:
            return lambda: nn.relu_bwd(dy, y, dx)
        return lambda: nn.add_act(dy, None, dx, 2, y)

    def cast(self, src, dst):
        return lambda: nn.cast(src, dst)

    def maxpool_fwd(self, x, y, code, k, s, p):
        return lambda: nn.maxpool_fwd(x, y, code, k, s, p)

    def maxpool_bwd(self, dy, code, dx, k, s, p):
        return lambda: nn.maxpool_bwd(dy, code, dx, k, s, p)

    def gap_fwd(self, x, y):
        return lambda: nn.gap_fwd(x, y)

    def gap_bwd(self, dy, dx, mask=None):
        return lambda: nn.gap_bwd(dy, dx, mask)

    def softmax_xent(self, logits, labels, dlogits, metrics, scale):
        return lambda: nn.softmax_xent(logits, labels, dlogits, metrics, scale)

    def sgd_apply(self, params, grads, n_params, lr):
        return lambda: nn.sgd_apply(params, grads, n_params, lr)
