"""Lowering of a layer stack to a fixed list of kernel launches (one training step).

The result is a ``StepProgram``: two flat lists of zero-argument callables — forward and
backward — over buffers that are allocated once.  Nothing in a step depends on host-side
values, which is what lets the CUDA backend capture the whole step in one graph and replay it
("hot path").

Conventions
-----------
* ``dy`` of an op is the gradient w.r.t. its *pre-activation* output.  The consumer that
  produces it applies the producer's ReLU mask in the same kernel when it can (GEMM dgrad
  epilogue, pool backward, fused head), so no stand-alone ReLU-backward pass runs over the
  big activations.
* Parameter gradients are *accumulated* into the gradient arena, which starts each step at
  zero (the fused all-reduce/SGD kernel re-zeroes it), so split-K and multi-CTA reductions can
  use atomics without a memset per step.
* Fusions (CUDA backend): first-layer Conv2D(C_in=1)+bias+ReLU+MaxPool in one kernel each
  way [K1+K2, K8+K9]; final Dense(C<=32)+softmax-xent+all gradients in one kernel [K4+K5+K6].
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
    first_param_layer: int = -1       # index of that layer (bucket boundary for overlapped all-reduce)
    description: list[str] = field(default_factory=list)
    keep: list = field(default_factory=list)        # buffers / plans kept alive

    def run(self):
        for f in self.forward:
            f()
        for f in self.backward:
            f()


def lower(layers, batch: int, be, layout: ArenaLayout, params, grads, x_in, labels,
          training: bool = True) -> StepProgram:
    """Build the step over ``x_in`` [B,H,W,C] / ``labels`` [B] for backend ``be``."""
    prog = StepProgram()
    metrics = grads[layout.n_params:layout.n_total]
    grad_scale = 1.0 / batch

    def pview(flat, layer_idx):
        sl = layout.layer_slots(layer_idx)
        k = layout.view(flat, sl[0])
        b = layout.view(flat, sl[1]) if len(sl) > 1 else None
        return k, b

    def _fuses_head(layer, is_last, k_in, units):
        return (training and is_last and layer.activation is None and layer.use_bias
                and getattr(be, "supports_conv_c1_pool", False) and be.supports_fused_head(k_in, units))

    pending_bias_grad = None
    # ---- forward walk --------------------------------------------------------------
    cur, d_cur = x_in, None             # activation and its gradient buffer (None: no grad needed)
    cur_post_relu = False               # True: gradient w.r.t. `cur` must be masked by cur > 0
    bwd_stages: list[list[Callable]] = []   # one list per op, emitted later in reverse
    i, n = 0, len(layers)
    head_done = False
    while i < n:
        layer = layers[i]
        nxt = layers[i + 1] if i + 1 < n else None
        is_last = i == n - 1
        if isinstance(layer, (L.InputLayer,)):
            i += 1
            continue
        if isinstance(layer, L.Flatten):
            cur = cur.view(batch, -1)
            d_cur = d_cur.view(batch, -1) if d_cur is not None else None
            prog.description.append(f"flatten -> {tuple(cur.shape)} (view)")
            i += 1
            continue
        if isinstance(layer, (L.Conv2D, L.Dense)):
            if prog.first_param_layer < 0:
                prog.first_param_layer = i
            elif prog.early_forward == 0:
                prog.early_forward = len(prog.forward)     # everything so far needs only layer 0's weights
        if isinstance(layer, L.Conv2D):
            w, b = pview(params, i)
            dw, db = pview(grads, i)
            ci = layer.input_shape[2]
            fuse_pool = (be.supports_conv_c1_pool and d_cur is None and ci == 1 and b is not None
                         and layer.kernel_size in ((3, 3), (5, 5)) and layer.strides == (1, 1)
                         and layer.padding == "valid" and layer.activation == "relu"
                         and isinstance(nxt, L.MaxPooling2D))
            if fuse_pool:
                ph, pw, co = nxt.output_shape
                y = be.alloc((batch, ph, pw, co))
                code = be.alloc((batch, ph, pw, co), torch.uint8)
                dy = be.alloc((batch, ph, pw, co)) if training else None
                prog.forward.append(be.conv_c1_pool_fwd(cur, w, b, y, code))
                if training:
                    bwd_stages.append([be.conv_c1_pool_bwd(cur, dy, code, dw, db)])
                prog.description.append(
                    f"conv{layer.kernel_size[0]}x{layer.kernel_size[1]}(C_in=1)+bias+relu+maxpool2 fused -> {tuple(y.shape)}")
                cur, d_cur, cur_post_relu = y, dy, False   # ReLU already folded into `code`
                i += 2
                continue
            oh, ow, co = layer.output_shape
            y = be.alloc((batch, oh, ow, co))
            dy = be.alloc((batch, oh, ow, co)) if training else None
            pad = layer.pad_amounts()
            relu = layer.activation == "relu"
            prog.forward.append(be.conv2d_fwd(cur, w, b, y, layer.strides, pad, relu))
            if training:
                stage = [be.conv2d_wgrad(cur, dy, dw, db, layer.strides, pad)]
                if d_cur is not None:
                    stage.append(be.conv2d_dgrad(dy, w, d_cur, layer.strides, pad,
                                                 mask=cur if cur_post_relu else None))
                bwd_stages.append(stage)
            prog.description.append(f"conv2d{'+relu' if relu else ''} -> {tuple(y.shape)} [{be.describe_conv(layer)}]")
            cur, d_cur, cur_post_relu = y, dy, relu
            i += 1
            continue
        if isinstance(layer, L.MaxPooling2D):
            ph, pw, c = layer.output_shape
            y = be.alloc((batch, ph, pw, c))
            code = be.alloc((batch, ph, pw, c), torch.uint8)
            dy = be.alloc((batch, ph, pw, c)) if training else None
            prog.forward.append(be.maxpool2_fwd(cur, y, code))
            if training and d_cur is not None:
                bwd_stages.append([be.maxpool2_bwd(dy, code, y, d_cur, cur_post_relu)])
            elif training:
                bwd_stages.append([])
            prog.description.append(f"maxpool2 -> {tuple(y.shape)}")
            # max(relu(x)) == relu(max(x)): the pooled map is still a post-ReLU map, but its
            # consumer must not mask again (the pool backward already applied the mask).
            cur, d_cur, cur_post_relu = y, dy, False
            i += 1
            continue
        if isinstance(layer, L.Dense):
            w, b = pview(params, i)
            dw, db = pview(grads, i)
            k_in, units = w.shape
            fuse_head = _fuses_head(layer, is_last, k_in, units)
            if fuse_head:
                logits = be.alloc((batch, units))
                prog.logits = logits
                stage = be.dense_xent_head(cur, w, b, labels, d_cur, dw, db, metrics, grad_scale,
                                           cur_post_relu, logits, dh_colsum=pending_bias_grad)
                pending_bias_grad = None
                # forward and backward of the head are one kernel; it runs at the fwd/bwd seam
                prog.forward.append(stage)
                bwd_stages.append([])
                prog.description.append(f"dense({units})+softmax_xent+grads fused head [K={k_in}]")
                head_done = True
                i += 1
                continue
            y = be.alloc((batch, units))
            dy = be.alloc((batch, units)) if training else None
            relu = layer.activation == "relu"
            prog.forward.append(be.dense_fwd(cur, w, b, y, relu))
            if training:
                # If the fused head consumes this layer's output it also produces this layer's
                # bias gradient (column sums of dy), so no separate reduction kernel is needed.
                head_takes_bias = (db is not None and isinstance(nxt, L.Dense) and i + 1 == n - 1
                                   and _fuses_head(nxt, True, units, nxt.units))
                if head_takes_bias:
                    pending_bias_grad = db
                stage = [be.dense_wgrad(cur, dy, dw, None if head_takes_bias else db)]
                if d_cur is not None:
                    stage.append(be.dense_dgrad(dy, w, d_cur, mask=cur if cur_post_relu else None))
                bwd_stages.append(stage)
            prog.description.append(f"dense({units}){'+relu' if relu else ''} [{be.describe_dense(batch, k_in, units)}]")
            cur, d_cur, cur_post_relu = y, dy, relu
            i += 1
            continue
        raise NotImplementedError(f"layer type {type(layer).__name__} is not supported by the engine")

    if not head_done:
        if cur.dim() != 2:
            raise ValueError("the model must end in a Dense layer producing logits")
        prog.logits = cur
        if training:
            if cur_post_relu:
                raise ValueError("the last layer must be linear (logits) for from_logits=True")
            prog.forward.append(be.softmax_xent(cur, labels, d_cur, metrics, grad_scale))
            prog.description.append("softmax_xent (+grad, +accuracy count)")

    for stage in reversed(bwd_stages):
        prog.backward.extend(stage)
    return prog
