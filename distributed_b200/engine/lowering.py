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
  BatchNorm+ReLU; residual add+ReLU, whose backward mask is applied by the kernel that writes the
  block's output gradient (the next block's gradient sum, or the pooling backward).
* Storage types: 4-D activations and their gradients use the backend's ``act_dtype`` (fp32, or
  bf16 under the ``mixed_bfloat16`` policy); 2-D (Dense) activations, parameters, parameter
  gradients and all per-channel statistics are fp32.  Flatten / GlobalAveragePooling2D are the
  boundaries where the conversion happens.
* BatchNormalization moving statistics are non-trainable arena slots; the forward kernel writes a
  pseudo-gradient for them so that the (replica-averaging) optimizer update performs
  ``m = momentum * m + (1 - momentum) * mean_over_replicas(batch statistic)``.
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
    early_backward: int = -1          # backward launches before the first parametrised layer's own backward stage
                                      # (everything after that layer is final there: bucket boundary); -1 = unknown
    first_param_layer: int = -1       # leaf index of that layer (bucket boundary for overlapped all-reduce)
    grad_final: list = field(default_factory=list)
    # grad_final: one (arena_begin, n_launches) pair per parametrised top-level layer, in forward order: once
    # the first n_launches backward launches have completed, every gradient in [arena_begin, n_params) is final
    # (the bucket schedule of the overlapped all-reduce is cut from this table, engine/trainer.py)
    description: list[str] = field(default_factory=list)
    keep: list = field(default_factory=list)        # buffers / plans kept alive

    def run(self):
        for f in self.forward:
            f()
        for f in self.backward:
            f()


def _side(f):
    """Tag a backward launch as independent of the rest of the backward chain (a weight gradient: it
    reads an activation and a gradient that are final, and writes only the gradient arena), so the
    trainer may put it on a concurrent branch of the step graph."""
    def run():
        f()
    run.side_ok = True
    run.writes_grads = True
    return run


def _grads(f):
    """Tag a backward launch that writes parameter gradients (the bucket schedule needs to know which launch
    is the LAST writer of a layer's gradients; data-gradient launches are not)."""
    return _tag(f, writes_grads=True)


def _tag(f, **attrs):
    """Wrap a launch and attach scheduling hints for the trainer's graph capture, keeping earlier ones:
    ``branch=True`` — the launch belongs to a residual block's shortcut branch, which depends only on the
    block's input (forward) / output gradient (backward) and may run on its own stream next to the main branch;
    ``join_branch=True`` — the launch is the first consumer of that branch's results."""
    def run():
        f()
    for k in ("side_ok", "branch", "join_branch", "writes_grads"):
        if getattr(f, k, False):
            setattr(run, k, True)
    for k, v in attrs.items():
        setattr(run, k, v)
    return run


@dataclass
class _State:
    cur: torch.Tensor                 # activation
    d_cur: torch.Tensor | None        # where the gradient w.r.t. `cur` must be written (None: not needed)
    post_relu: bool = False           # True: that gradient must be masked by cur > 0 by whoever writes it
    addend: torch.Tensor | None = None  # the writer must ADD this tensor to the gradient (residual sum)


class _Lowerer:
    def __init__(self, batch, be, layout, params, grads, labels, training, lr=0.0):
        self.batch, self.be, self.layout = batch, be, layout
        self.params, self.grads, self.labels, self.training = params, grads, labels, training
        self.inv_lr = 1.0 / lr if lr else 0.0
        self.act_dtype = getattr(be, "act_dtype", torch.float32)
        self.prog = StepProgram()
        self.metrics = grads[layout.n_params:layout.n_total]
        self.grad_scale = 1.0 / batch
        self.stages: list[list[Callable]] = []      # backward stages in forward order
        self.pending_bias_grad = None
        self.head_done = False
        self.fused_tail = False
        self.first_param_stage = -1
        self.top_marks: list = []         # (top-level layer, index of its first backward stage)

    # -- helpers ---------------------------------------------------------------------
    def pview(self, flat, layer):
        return [self.layout.view(flat, s) for s in self.layout.layer_slots(layer)]

    def kernel_bias(self, flat, layer):
        v = self.pview(flat, layer)
        return v[0], (v[1] if len(v) > 1 else None)

    def fuses_head(self, layer, is_last, k_in, units):
        return (self.training and is_last and layer.activation is None and layer.use_bias
                and getattr(self.be, "supports_conv_c1_pool", False) and self.be.supports_fused_head(k_in, units))

    def note_param_layer(self, layer, top_level=True):
        prog = self.prog
        if prog.first_param_layer < 0:
            prog.first_param_layer = self.layout.leaf_index(layer)
            self.first_param_stage = len(self.stages) if top_level else -1
        elif prog.early_forward == 0:
            prog.early_forward = len(prog.forward)     # everything so far needs only the first layer's weights

    def alloc_grad(self, shape, dtype=torch.float32):
        return self.be.alloc(shape, dtype) if self.training else None

    def act(self, shape):
        """A 4-D activation buffer in the backend's activation storage type."""
        return self.be.alloc(shape, self.act_dtype)

    def act_grad(self, shape, final_dy=None):
        """Gradient buffer of a 4-D activation; ``final_dy`` (the enclosing residual block's output
        gradient) is used instead when the layer is the last one of a branch."""
        if not self.training:
            return None
        return final_dy if final_dy is not None else self.be.alloc(shape, self.act_dtype)

    def relu_mask(self, d_buf, act):
        """In-place d_buf *= (act > 0) for consumers that cannot fuse the mask."""
        return self.be.relu_bwd(d_buf, act, d_buf)

    # -- a list of layers --------------------------------------------------------------
    def sequence(self, layers, st: _State, top_level: bool, final_dy=None, final_fuse=None) -> _State:
        """Lower ``layers`` starting from state ``st``.  ``final_dy``: gradient buffer the LAST layer
        must use for its output (a residual branch reads the block's output gradient directly).
        ``final_fuse`` = (residual, relu, out): if the last layer is a BatchNormalization it computes
        out = act(bn(x) + residual) itself (the tail of a residual block); sets ``self.fused_tail``."""
        be, batch, prog, training = self.be, self.batch, self.prog, self.training
        i, n = 0, len(layers)
        while i < n:
            layer = layers[i]
            nxt = layers[i + 1] if i + 1 < n else None
            is_last = top_level and i == n - 1
            fdy = final_dy if i == n - 1 else None
            if isinstance(layer, L.InputLayer):
                i += 1
                continue
            if isinstance(layer, L.Flatten):
                st = self.flatten(st)
                i += 1
                continue
            if isinstance(layer, (L.Conv2D, L.Dense, L.BatchNormalization)):
                self.note_param_layer(layer, top_level)
            if top_level:
                self.top_marks.append((layer, len(self.stages)))

            if isinstance(layer, L.Conv2D):
                w, b = self.kernel_bias(self.params, layer)
                dw, db = self.kernel_bias(self.grads, layer)
                ci = layer.input_shape[2]
                fuse_pool = (be.supports_conv_c1_pool and st.d_cur is None and ci == 1 and b is not None
                             and self.act_dtype == torch.float32
                             and layer.kernel_size in ((3, 3), (5, 5)) and layer.strides == (1, 1)
                             and layer.padding == "valid" and layer.activation == "relu"
                             and isinstance(nxt, L.MaxPooling2D) and nxt.is_default_2x2
                             and not (final_dy is not None and i + 1 == n - 1))
                if fuse_pool:
                    ph, pw, co = nxt.output_shape
                    y = be.alloc((batch, ph, pw, co))
                    code = be.alloc((batch, ph, pw, co), torch.uint8)
                    dy = self.alloc_grad((batch, ph, pw, co))
                    prog.forward.append(be.conv_c1_pool_fwd(st.cur, w, b, y, code))
                    if training:
                        self.stages.append([_grads(be.conv_c1_pool_bwd(st.cur, dy, code, dw, db))])
                    prog.description.append(
                        f"conv{layer.kernel_size[0]}x{layer.kernel_size[1]}(C_in=1)+bias+relu+maxpool2 fused -> {tuple(y.shape)}")
                    st = _State(y, dy, False)       # ReLU already folded into `code`
                    i += 2
                    continue
                oh, ow, co = layer.output_shape
                y = self.act((batch, oh, ow, co))
                dy = self.act_grad((batch, oh, ow, co), fdy)
                pad = layer.pad_amounts()
                relu = layer.activation == "relu"
                prog.forward.append(be.conv2d_fwd(st.cur, w, b, y, layer.strides, pad, relu))
                if training:
                    stage = [_side(be.conv2d_wgrad(st.cur, dy, dw, db, layer.strides, pad))]
                    if st.d_cur is not None:
                        extra = {"addend": st.addend} if st.addend is not None else {}
                        stage.append(be.conv2d_dgrad(dy, w, st.d_cur, layer.strides, pad,
                                                     mask=st.cur if st.post_relu else None, **extra))
                    self.stages.append(stage)
                prog.description.append(
                    f"conv2d{layer.kernel_size}{'/s' + str(layer.strides[0]) if layer.strides != (1, 1) else ''}"
                    f"{'+relu' if relu else ''} -> {tuple(y.shape)} [{be.describe_conv(layer)}]")
                st = _State(y, dy, relu)
                i += 1
                continue

            if isinstance(layer, L.MaxPooling2D):
                ph, pw, c = layer.output_shape
                y = self.act((batch, ph, pw, c))
                code = be.alloc((batch, ph, pw, c), torch.uint8)
                dy = self.act_grad((batch, ph, pw, c), fdy)
                if layer.is_default_2x2 and self.act_dtype == torch.float32:
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
                                stage.append(self.relu_mask(st.d_cur, st.cur))
                        self.stages.append(stage)
                prog.description.append(f"maxpool{layer.pool_size[0]}/s{layer.strides[0]} -> {tuple(y.shape)}")
                # max(relu(x)) == relu(max(x)): the pool backward already applied the producer's mask
                st = _State(y, dy, False)
                i += 1
                continue

            if isinstance(layer, L.BatchNormalization):
                st = self.batchnorm(layer, st, fdy, final_fuse if i == n - 1 else None)
                i += 1
                continue

            if isinstance(layer, L.GlobalAveragePooling2D):
                c = layer.input_shape[-1]
                y = be.alloc((batch, c))                    # fp32 from here on
                dy = self.alloc_grad((batch, c)) if fdy is None else fdy
                prog.forward.append(be.gap_fwd(st.cur, y))
                if training:
                    self.stages.append([be.gap_bwd(dy, st.d_cur, st.cur if st.post_relu else None)]
                                       if st.d_cur is not None else [])
                prog.description.append(f"global_average_pool -> {tuple(y.shape)}")
                st = _State(y, dy, False)
                i += 1
                continue

            if isinstance(layer, L.Residual):
                st = self.residual(layer, st, fdy)
                i += 1
                continue

            if isinstance(layer, L.Activation):
                # stand-alone activation (one that could not be folded into its producer, e.g. after a pooling
                # layer or a residual block): y = relu(x); the consumer masks dy with y > 0 (post_relu), so the
                # backward pass is a copy
                if layer.activation is None or st.post_relu:
                    i += 1                              # identity / relu(relu(x)): nothing to do
                    continue
                if layer.activation != "relu":
                    raise NotImplementedError(f"activation {layer.activation!r} is not supported")
                if st.cur.dim() != 4:
                    raise NotImplementedError("a stand-alone ReLU is supported on 4-D activations only; "
                                              "pass activation='relu' to the Dense layer instead")
                shape = tuple(st.cur.shape)
                y = be.alloc(shape, st.cur.dtype)
                dy = (fdy if fdy is not None else be.alloc(shape, st.cur.dtype)) if training else None
                prog.forward.append(be.add_act(st.cur, None, y, 1))
                if training:
                    self.stages.append([be.add_act(dy, None, st.d_cur, 0)] if st.d_cur is not None else [])
                prog.description.append(f"relu -> {shape}")
                st = _State(y, dy, True)
                i += 1
                continue

            if isinstance(layer, L.Dense):
                if st.cur.dim() != 2 or st.cur.dtype != torch.float32:
                    raise ValueError("Dense expects a flattened fp32 input: put Flatten or "
                                     "GlobalAveragePooling2D before it")
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
                dy = self.alloc_grad((batch, units)) if fdy is None else fdy
                relu = layer.activation == "relu"
                prog.forward.append(be.dense_fwd(st.cur, w, b, y, relu))
                if training:
                    # If the fused head consumes this layer's output it also produces this layer's
                    # bias gradient (column sums of dy), so no separate reduction kernel is needed.
                    head_takes_bias = (db is not None and top_level and isinstance(nxt, L.Dense) and i + 1 == n - 1
                                       and self.fuses_head(nxt, True, units, nxt.units))
                    if head_takes_bias:
                        self.pending_bias_grad = db
                    stage = [_side(be.dense_wgrad(st.cur, dy, dw, None if head_takes_bias else db))]
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

    # -- Flatten: a view in fp32; the fp32 <-> bf16 boundary under the mixed policy -----------
    def flatten(self, st: _State) -> _State:
        be, batch, prog = self.be, self.batch, self.prog
        if st.cur.dtype == torch.float32:
            prog.description.append(f"flatten -> {tuple(st.cur.view(batch, -1).shape)} (view)")
            return _State(st.cur.view(batch, -1), st.d_cur.view(batch, -1) if st.d_cur is not None else None,
                          st.post_relu)
        flat = st.cur.view(batch, -1)
        y = be.alloc(tuple(flat.shape))
        dy = self.alloc_grad(tuple(flat.shape))
        prog.forward.append(be.cast(flat, y))
        if self.training:
            stage = []
            if st.d_cur is not None:
                stage.append(be.cast(dy, st.d_cur.view(batch, -1)))
                if st.post_relu:
                    stage.append(self.relu_mask(st.d_cur, st.cur))
            self.stages.append(stage)
        prog.description.append(f"flatten -> {tuple(y.shape)} (bf16 -> fp32)")
        return _State(y, dy, False)

    # -- BatchNormalization (+ReLU) ----------------------------------------------------------
    def batchnorm(self, layer, st: _State, fdy, fuse=None) -> _State:
        be, prog, training = self.be, self.prog, self.training
        gamma, beta, mmean, mvar = self.pview(self.params, layer)
        dgamma, dbeta, g_mmean, g_mvar = self.pview(self.grads, layer)
        c = layer.input_shape[-1]
        relu = layer.activation == "relu"
        shape = tuple(st.cur.shape)
        extra = {}
        if fuse is not None and not relu and getattr(be, "supports_bn_residual", False):
            res, relu, y = fuse                       # y = act(bn(x) + residual), written to the block output
            extra = {"residual": res.view(-1, c)}
            self.fused_tail = True
        else:
            y = be.alloc(shape, st.cur.dtype)
        dy = (fdy if fdy is not None else be.alloc(shape, st.cur.dtype)) if training else None
        x2d, y2d = st.cur.view(-1, c), y.view(-1, c)
        if training:
            stats, saved, red = be.alloc((2 * c,)), be.alloc((2 * c,)), be.alloc((2 * c,))
            dx2d = st.d_cur.view(-1, c) if st.d_cur is not None else None
            fwd, bwd = be.batchnorm(x2d, y2d, gamma, beta, eps=layer.epsilon, momentum=layer.momentum, relu=relu,
                                    stats=stats, saved=saved, red=red, moving=(mmean, mvar),
                                    moving_grads=(g_mmean, g_mvar), inv_lr=self.inv_lr, inference=False,
                                    dy2d=dy.view(-1, c), dx2d=dx2d, dgamma=dgamma, dbeta=dbeta, **extra)
            stage = [_grads(bwd)]
            if st.d_cur is not None and st.post_relu:
                stage.append(self.relu_mask(st.d_cur, st.cur))
            self.stages.append(stage)
        else:
            fwd, _ = be.batchnorm(x2d, y2d, gamma, beta, eps=layer.epsilon, momentum=layer.momentum, relu=relu,
                                  moving=(mmean, mvar), inference=True, **extra)
        prog.forward.append(fwd)
        prog.description.append(f"batchnorm{'+residual' if extra else ''}{'+relu' if relu else ''} [C={c}]"
                                f"{'' if training else ' (moving statistics)'}")
        # own ReLU: masks internally from its output.  Fused block tail: the block's ReLU is the CONSUMER's
        # business (post_relu), exactly as for an unfused residual block.
        return _State(y, dy, relu if extra else False)

    # -- relu(main(x) + shortcut(x)) ----------------------------------------------------
    def residual(self, layer, st: _State, fdy=None) -> _State:
        """Forward: y = act(main(x) + shortcut(x)).  Backward: the block's output gradient dy arrives
        already masked by (y > 0) — ``post_relu`` on the returned state makes its writer do that — and
        is read directly by the last layer of BOTH branches, so the block costs no ReLU-backward pass
        and no copy; d_x = d_main + d_shortcut is one (masked) add."""
        be, training, prog = self.be, self.training, self.prog
        x, shape = st.cur, tuple(st.cur.shape)
        need_dx = training and st.d_cur is not None
        relu = layer.activation == "relu"
        d_sc_in = be.alloc(shape, x.dtype) if (need_dx and layer.shortcut) else None
        out_shape = (self.batch, *layer.output_shape)
        y = be.alloc(out_shape, x.dtype)
        dy = (fdy if fdy is not None else be.alloc(out_shape, x.dtype)) if training else None
        sum_stage: list[Callable] = []
        self.stages.append(sum_stage)               # runs LAST in backward: d_x = d_main_in + d_shortcut_in
        prog.description.append(f"residual block {layer.name}:")
        # d_x = d_main + d_shortcut (masked by x > 0 when x is a ReLU output).  When the main branch
        # starts with a convolution whose data gradient runs on the tensor cores, that kernel adds the
        # shortcut's gradient (and masks) in its epilogue and writes d_x itself; otherwise one add kernel.
        other = (d_sc_in if layer.shortcut else dy) if need_dx else None
        first = layer.main[0] if layer.main else None
        fuse_sum = (need_dx and isinstance(first, L.Conv2D) and hasattr(be, "dgrad_addend_ok")
                    and be.dgrad_addend_ok(shape, (*first.kernel_size, shape[3], first.filters), first.strides,
                                           first.pad_amounts(), first.output_shape[:2], x.dtype))
        if fuse_sum:
            main_in = _State(x, st.d_cur, st.post_relu, addend=other)
        else:
            d_main_in = be.alloc(shape, x.dtype) if need_dx else None
            main_in = _State(x, d_main_in, False)
        # Forward order: shortcut, then main — so that a BatchNorm ending the main branch can add the
        # shortcut and apply the block's ReLU itself.  Backward order is the reverse of the order of
        # self.stages, and must be: shortcut, then main (whose first convolution may consume the shortcut's
        # input gradient), then the sum — hence the stage lists are swapped after lowering.
        i0, f0 = len(self.stages), len(prog.forward)
        s = (self.sequence(layer.shortcut, _State(x, d_sc_in, False), top_level=False, final_dy=dy)
             if layer.shortcut else None)
        i1, f1 = len(self.stages), len(prog.forward)
        # the shortcut branch is independent of the main branch: let the trainer run it on its own stream
        prog.forward[f0:f1] = [_tag(f, branch=True) for f in prog.forward[f0:f1]]
        for k in range(i0, i1):
            self.stages[k][:] = [_tag(f, branch=True) for f in self.stages[k]]
        if s is not None and s.post_relu:
            raise ValueError("the last layer of a residual branch must not carry an activation")
        sc_out = s.cur if s is not None else x
        self.fused_tail = False
        m = self.sequence(layer.main, main_in, top_level=False, final_dy=dy, final_fuse=(sc_out, relu, y))
        fused_tail, self.fused_tail = self.fused_tail, False
        if f1 > f0:
            # first consumers of the shortcut: forward, the launch that adds it (last of the main branch when the
            # BatchNorm tail is fused, else the add below); backward, whichever launch folds its input gradient in
            # (the last launch of the main branch's first stage = its first convolution's dgrad, else the sum stage)
            if fused_tail:
                prog.forward[-1] = _tag(prog.forward[-1], join_branch=True)
            if fuse_sum and i1 < len(self.stages) and self.stages[i1]:
                self.stages[i1][-1] = _tag(self.stages[i1][-1], join_branch=True)
        self.stages[i0:] = self.stages[i1:] + self.stages[i0:i1]
        if tuple(m.cur.shape) != out_shape or (s is not None and tuple(s.cur.shape) != out_shape):
            raise ValueError(f"residual block {layer.name}: branch output shapes differ")
        if not fused_tail:
            if m.post_relu:
                raise ValueError("the last layer of a residual branch must not carry an activation")
            prog.forward.append(_tag(be.add_act(m.cur, sc_out, y, 1 if relu else 0), join_branch=f1 > f0))
        if need_dx and not fuse_sum:
            if st.post_relu:
                sum_stage.append(_tag(be.add_act(d_main_in, other, st.d_cur, 2, mask=x), join_branch=f1 > f0))
            else:
                sum_stage.append(_tag(be.add_act(d_main_in, other, st.d_cur, 0), join_branch=f1 > f0))
        if fuse_sum:
            prog.description.append("  gradient sum fused into the first convolution's data-gradient epilogue")
        prog.description.append(f"  add{'+relu' if relu else ''} -> {out_shape}"
                                f"{' (inside the last batchnorm)' if fused_tail else ''}")
        return _State(y, dy, relu)


def lower(layers, batch: int, be, layout: ArenaLayout, params, grads, x_in, labels,
          training: bool = True, lr: float = 0.0) -> StepProgram:
    """Build the step over ``x_in`` [B,H,W,C] / ``labels`` [B] for backend ``be``.  ``lr`` is the
    optimizer's learning rate (scales the BatchNorm moving-statistics pseudo-gradients)."""
    lw = _Lowerer(batch, be, layout, params, grads, labels, training, lr)
    begin = be.begin_step(params) if hasattr(be, "begin_step") else None
    if begin is not None:
        lw.prog.forward.append(begin)
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
    if training and lw.first_param_stage >= 0:
        # stages after the first parametrised layer's run first in backward
        prog.early_backward = sum(len(stage) for stage in lw.stages[lw.first_param_stage + 1:])
    if training:
        from ..keras.layers import flatten_layers
        for layer, stage0 in lw.top_marks:
            offs = [s.offset for leaf in flatten_layers([layer]) for s in layout.layer_slots(leaf)]
            if not offs:
                continue
            n_prefix = sum(len(stage) for stage in lw.stages[stage0:])       # launches of layers >= this one
            last = max((k for k in range(n_prefix) if getattr(prog.backward[k], "writes_grads", False)), default=-1)
            prog.grad_final.append((min(offs), last + 1))
    if begin is not None:
        prog.early_forward = 0       # the step starts by re-reading every weight: nothing to overlap
    return prog
