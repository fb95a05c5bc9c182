"""Lowering of a layer stack to a fixed list of kernel launches (one training step).

The result is a ``StepProgram``: two flat lists of zero-argument callables — forward and
backward — over buffers that are allocated once.  Nothing in a step depends on host-side
values, which is what lets the CUDA backend capture the whole step in one graph and replay it
("hot path").

Conventions
-----------
* ``dy`` of an op is the gradient w.r.t. its output as its consumer sees it.  For Conv2D / Dense
  with a fused ReLU that means the *pre-activation* gradient: the consumer that produces it
  applies the ReLU mask in the same kernel when it can (GEMM dgrad epilogue, pool backward,
  fused head), so no stand-alone ReLU-backward pass runs over the big activations.  Ops that own
  their ReLU (BatchNorm+ReLU, residual add+ReLU) mask internally from their saved output.
* Parameter gradients are *accumulated* into the gradient arena, which starts each step at
  zero (the fused all-reduce/SGD kernel re-zeroes it), so split-K and multi-CTA reductions can
  use atomics without a memset per step.
* Fusions (CUDA backend): first-layer Conv2D(C_in=1)+bias+ReLU+MaxPool in one kernel each
  way [K1+K2, K8+K9]; final Dense(C<=32)+softmax-xent+all gradients in one kernel [K4+K5+K6];
  BatchNorm+ReLU; residual add+ReLU.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable

import torch

from ..keras import layers as L
from .arena import ArenaLayout


@dataclass
class StepProgram:
    forward: list[Callable[[], None]] = field(default_factory=list)
    backward: list[Callable[[], None]] = field(default_factory=list)
    logits: torch.Tensor | None = None
    early_forward: int = 0            # forward launches that need only the FIRST parametrised layer
    first_param_layer: int = -1       # leaf index of that layer (bucket boundary for overlapped all-reduce)
    description: list[str] = field(default_factory=list)
    keep: list = field(default_factory=list)        # buffers / plans kept alive

    def run(self):
        for f in self.forward:
            f()
        for f in self.backward:
            f()


@dataclass
class _State:
    cur: torch.Tensor                 # activation
    d_cur: torch.Tensor | None        # where the gradient w.r.t. `cur` must be written (None: not needed)
    post_relu: bool = False           # True: that gradient must be masked by cur > 0 by whoever writes it


class _Lowerer:
    def __init__(self, batch, be, layout, params, grads, labels, training):
        self.batch, self.be, self.layout = batch, be, layout
        self.params, self.grads, self.labels, self.training = params, grads, labels, training
        self.prog = StepProgram()
        self.metrics = grads[layout.n_params:layout.n_total]
        self.grad_scale = 1.0 / batch
        self.stages: list[list[Callable]] = []      # backward stages in forward order
        self.pending_bias_grad = None
        self.head_done = False

    # -- helpers ---------------------------------------------------------------------
    def pview(self, flat, layer):
        return [self.layout.view(flat, s) for s in self.layout.layer_slots(layer)]

    def kernel_bias(self, flat, layer):
        v = self.pview(flat, layer)
        return v[0], (v[1] if len(v) > 1 else None)

    def fuses_head(self, layer, is_last, k_in, units):
        return (self.training and is_last and layer.activation is None and layer.use_bias
                and getattr(self.be, "supports_conv_c1_pool", False) and self.be.supports_fused_head(k_in, units))

    def note_param_layer(self, layer):
        prog = self.prog
        if prog.first_param_layer < 0:
            prog.first_param_layer = self.layout.leaf_index(layer)
        elif prog.early_forward == 0:
            prog.early_forward = len(prog.forward)     # everything so far needs only the first layer's weights

    def alloc_grad(self, shape):
        return self.be.alloc(shape) if self.training else None

    # -- a list of layers --------------------------------------------------------------
    def sequence(self, layers, st: _State, top_level: bool) -> _State:
        be, batch, prog, training = self.be, self.batch, self.prog, self.training
        i, n = 0, len(layers)
        while i < n:
            layer = layers[i]
            nxt = layers[i + 1] if i + 1 < n else None
            is_last = top_level and i == n - 1
            if isinstance(layer, L.InputLayer):
                i += 1
                continue
            if isinstance(layer, L.Flatten):
                st = _State(st.cur.view(batch, -1), st.d_cur.view(batch, -1) if st.d_cur is not None else None,
                            st.post_relu)
                prog.description.append(f"flatten -> {tuple(st.cur.shape)} (view)")
                i += 1
                continue
            if isinstance(layer, (L.Conv2D, L.Dense, L.BatchNormalization)):
                self.note_param_layer(layer)

            if isinstance(layer, L.Conv2D):
                w, b = self.kernel_bias(self.params, layer)
                dw, db = self.kernel_bias(self.grads, layer)
                ci = layer.input_shape[2]
                fuse_pool = (be.supports_conv_c1_pool and st.d_cur is None and ci == 1 and b is not None
                             and layer.kernel_size in ((3, 3), (5, 5)) and layer.strides == (1, 1)
                             and layer.padding == "valid" and layer.activation == "relu"
                             and isinstance(nxt, L.MaxPooling2D) and nxt.is_default_2x2)
                if fuse_pool:
                    ph, pw, co = nxt.output_shape
                    y = be.alloc((batch, ph, pw, co))
                    code = be.alloc((batch, ph, pw, co), torch.uint8)
                    dy = self.alloc_grad((batch, ph, pw, co))
                    prog.forward.append(be.conv_c1_pool_fwd(st.cur, w, b, y, code))
                    if training:
                        self.stages.append([be.conv_c1_pool_bwd(st.cur, dy, code, dw, db)])
                    prog.description.append(
                        f"conv{layer.kernel_size[0]}x{layer.kernel_size[1]}(C_in=1)+bias+relu+maxpool2 fused -> {tuple(y.shape)}")
                    st = _State(y, dy, False)       # ReLU already folded into `code`
                    i += 2
                    continue
                oh, ow, co = layer.output_shape
                y = be.alloc((batch, oh, ow, co))
                dy = self.alloc_grad((batch, oh, ow, co))
                pad = layer.pad_amounts()
                relu = layer.activation == "relu"
                prog.forward.append(be.conv2d_fwd(st.cur, w, b, y, layer.strides, pad, relu))
                if training:
                    stage = [be.conv2d_wgrad(st.cur, dy, dw, db, layer.strides, pad)]
                    if st.d_cur is not None:
                        stage.append(be.conv2d_dgrad(dy, w, st.d_cur, layer.strides, pad,
                                                     mask=st.cur if st.post_relu else None))
                    self.stages.append(stage)
                prog.description.append(
                    f"conv2d{layer.kernel_size}{'/s' + str(layer.strides[0]) if layer.strides != (1, 1) else ''}"
                    f"{'+relu' if relu else ''} -> {tuple(y.shape)} [{be.describe_conv(layer)}]")
                st = _State(y, dy, relu)
                i += 1
                continue

            if isinstance(layer, L.MaxPooling2D):
                ph, pw, c = layer.output_shape
                y = be.alloc((batch, ph, pw, c))
                code = be.alloc((batch, ph, pw, c), torch.uint8)
                dy = self.alloc_grad((batch, ph, pw, c))
                if layer.is_default_2x2:
                    prog.forward.append(be.maxpool2_fwd(st.cur, y, code))
                    if training:
                        self.stages.append([be.maxpool2_bwd(dy, code, y, st.d_cur, st.post_relu)]
                                           if st.d_cur is not None else [])
                else:
                    k, s, p = layer.pool_size[0], layer.strides[0], layer.pad_amount()
                    prog.forward.append(be.maxpool_fwd(st.cur, y, code, k, s, p))
                    if training:
                        stage = []
                        if st.d_cur is not None:
                            stage.append(be.maxpool_bwd(dy, code, st.d_cur, k, s, p))
                            if st.post_relu:
                                stage.append(be.relu_bwd(st.d_cur, st.cur, st.d_cur))
                        self.stages.append(stage)
                prog.description.append(f"maxpool{layer.pool_size[0]}/s{layer.strides[0]} -> {tuple(y.shape)}")
                # max(relu(x)) == relu(max(x)): the pool backward already applied the producer's mask
                st = _State(y, dy, False)
                i += 1
                continue

            if isinstance(layer, L.BatchNormalization):
                gamma, beta = self.pview(self.params, layer)
                dgamma, dbeta = self.pview(self.grads, layer)
                c = layer.input_shape[-1]
                relu = layer.activation == "relu"
                y = be.alloc(tuple(st.cur.shape))
                dy = self.alloc_grad(tuple(st.cur.shape))
                stats, saved, red = be.alloc((2 * c,)), be.alloc((2 * c,)), be.alloc((2 * c,))
                if getattr(layer, "_moving", None) is None or layer._moving.device != st.cur.device:
                    moving = be.alloc((2 * c,))
                    moving[c:] = 1.0
                    layer._moving = moving          # moving_mean | moving_variance, exported with the weights
                moving = layer._moving
                x2d, y2d = st.cur.view(-1, c), y.view(-1, c)
                prog.forward.append(be.bn_fwd(x2d, gamma, beta, stats, saved, moving if training else None, y2d,
                                              layer.epsilon, layer.momentum, relu))
                if training:
                    dx2d = st.d_cur.view(-1, c) if st.d_cur is not None else None
                    stage = [be.bn_bwd(x2d, y2d, dy.view(-1, c), gamma, saved, red, dx2d, dgamma, dbeta, relu)]
                    if st.d_cur is not None and st.post_relu:
                        stage.append(be.relu_bwd(st.d_cur, st.cur, st.d_cur))
                    self.stages.append(stage)
                prog.description.append(f"batchnorm{'+relu' if relu else ''} [C={c}]")
                st = _State(y, dy, False)          # owns its ReLU mask
                i += 1
                continue

            if isinstance(layer, L.GlobalAveragePooling2D):
                c = layer.input_shape[-1]
                y = be.alloc((batch, c))
                dy = self.alloc_grad((batch, c))
                prog.forward.append(be.gap_fwd(st.cur, y))
                if training:
                    stage = []
                    if st.d_cur is not None:
                        stage.append(be.gap_bwd(dy, st.d_cur))
                        if st.post_relu:
                            stage.append(be.relu_bwd(st.d_cur, st.cur, st.d_cur))
                    self.stages.append(stage)
                prog.description.append(f"global_average_pool -> {tuple(y.shape)}")
                st = _State(y, dy, False)
                i += 1
                continue

            if isinstance(layer, L.Residual):
                st = self.residual(layer, st)
                i += 1
                continue

            if isinstance(layer, L.Dense):
                w, b = self.kernel_bias(self.params, layer)
                dw, db = self.kernel_bias(self.grads, layer)
                k_in, units = w.shape
                if self.fuses_head(layer, is_last, k_in, units):
                    logits = be.alloc((batch, units))
                    prog.logits = logits
                    stage = be.dense_xent_head(st.cur, w, b, self.labels, st.d_cur, dw, db, self.metrics,
                                               self.grad_scale, st.post_relu, logits,
                                               dh_colsum=self.pending_bias_grad)
                    self.pending_bias_grad = None
                    prog.forward.append(stage)      # forward and backward of the head are one kernel
                    self.stages.append([])
                    prog.description.append(f"dense({units})+softmax_xent+grads fused head [K={k_in}]")
                    self.head_done = True
                    i += 1
                    continue
                y = be.alloc((batch, units))
                dy = self.alloc_grad((batch, units))
                relu = layer.activation == "relu"
                prog.forward.append(be.dense_fwd(st.cur, w, b, y, relu))
                if training:
                    # If the fused head consumes this layer's output it also produces this layer's
                    # bias gradient (column sums of dy), so no separate reduction kernel is needed.
                    head_takes_bias = (db is not None and top_level and isinstance(nxt, L.Dense) and i + 1 == n - 1
                                       and self.fuses_head(nxt, True, units, nxt.units))
                    if head_takes_bias:
                        self.pending_bias_grad = db
                    stage = [be.dense_wgrad(st.cur, dy, dw, None if head_takes_bias else db)]
                    if st.d_cur is not None:
                        stage.append(be.dense_dgrad(dy, w, st.d_cur, mask=st.cur if st.post_relu else None))
                    self.stages.append(stage)
                prog.description.append(
                    f"dense({units}){'+relu' if relu else ''} [{be.describe_dense(batch, k_in, units)}]")
                st = _State(y, dy, relu)
                i += 1
                continue
            raise NotImplementedError(f"layer type {type(layer).__name__} is not supported by the engine")
        return st

    # -- relu(main(x) + shortcut(x)) ----------------------------------------------------
    def residual(self, layer, st: _State) -> _State:
        be, training, prog = self.be, self.training, self.prog
        x, shape = st.cur, tuple(st.cur.shape)
        need_dx = training and st.d_cur is not None
        d_main_in = be.alloc(shape) if need_dx else None
        d_sc_in = be.alloc(shape) if (need_dx and layer.shortcut) else None
        sum_stage: list[Callable] = []
        self.stages.append(sum_stage)               # runs LAST in backward: d_x = d_main_in + d_shortcut_in
        prog.description.append(f"residual block {layer.name}:")
        m = self.sequence(layer.main, _State(x, d_main_in, False), top_level=False)
        s = self.sequence(layer.shortcut, _State(x, d_sc_in, False), top_level=False) if layer.shortcut else None
        if m.post_relu or (s is not None and s.post_relu):
            raise ValueError("the last layer of a residual branch must not carry an activation")
        out_shape = tuple(m.cur.shape)
        y = be.alloc(out_shape)
        dy = self.alloc_grad(out_shape)
        relu = layer.activation == "relu"
        sc_out = s.cur if s is not None else x
        prog.forward.append(be.add_act(m.cur, sc_out, y, relu))
        if training:
            # dz = dy * (y > 0) is the gradient of BOTH branches' outputs: computed once into the main
            # branch's buffer and copied for the shortcut branch.
            dz = m.d_cur
            stage = [be.relu_bwd(dy, y, dz)] if relu else [be.copy(dy, dz)]
            if s is not None:
                stage.append(be.copy(dz, s.d_cur))
            self.stages.append(stage)
            if need_dx:
                other = d_sc_in if s is not None else dz
                sum_stage.append(be.add_act(d_main_in, other, st.d_cur, False))
                if st.post_relu:
                    sum_stage.append(be.relu_bwd(st.d_cur, x, st.d_cur))
        prog.description.append(f"  add{'+relu' if relu else ''} -> {out_shape}")
        return _State(y, dy, False)


def lower(layers, batch: int, be, layout: ArenaLayout, params, grads, x_in, labels,
          training: bool = True) -> StepProgram:
    """Build the step over ``x_in`` [B,H,W,C] / ``labels`` [B] for backend ``be``."""
    lw = _Lowerer(batch, be, layout, params, grads, labels, training)
    st = lw.sequence(layers, _State(x_in, None, False), top_level=True)
    prog = lw.prog
    if not lw.head_done:
        if st.cur.dim() != 2:
            raise ValueError("the model must end in a Dense layer producing logits")
        prog.logits = st.cur
        if training:
            if st.post_relu:
                raise ValueError("the last layer must be linear (logits) for from_logits=True")
            prog.forward.append(be.softmax_xent(st.cur, labels, st.d_cur, lw.metrics, lw.grad_scale))
            prog.description.append("softmax_xent (+grad, +accuracy count)")
    for stage in reversed(lw.stages):
        prog.backward.extend(stage)
    return prog
