"""Trainer: owns the arenas, the lowered step and the step loop of one compiled model.

CUDA: the step (forward, loss, backward, fused all-reduce + SGD) is captured into CUDA graphs
by the native executor (csrc/runtime/executor.cpp) and replayed from a C++ loop that also
prefetches the next batch on a copy stream and streams the metric sums back to pinned host
memory — no Python, no host sync inside the loop.
CPU: the same lowered program runs op by op through the C++ CPU backend.
"""
from __future__ import annotations

import ctypes as C
import time

import numpy as np
import torch

from .arena import ArenaLayout, TAIL
from .lowering import lower


def _declare_exec(lib):
    if getattr(lib, "_exec_declared", False):
        return
    vp, i, i64, sz = C.c_void_p, C.c_int, C.c_int64, C.c_size_t
    lib.b2d_exec_create.argtypes = [C.POINTER(vp)]
    lib.b2d_exec_destroy.argtypes = [vp]
    lib.b2d_exec_destroy.restype = None
    lib.b2d_exec_stream.argtypes = [vp]
    lib.b2d_exec_stream.restype = vp
    lib.b2d_exec_side_stream.argtypes = [vp]
    lib.b2d_exec_side_stream.restype = vp
    lib.b2d_exec_fork.argtypes = [vp]
    lib.b2d_exec_join.argtypes = [vp]
    lib.b2d_exec_set_io.argtypes = [vp, i, vp, sz, vp, sz]
    lib.b2d_exec_set_metrics.argtypes = [vp, vp, sz, vp, i]
    lib.b2d_exec_begin_capture.argtypes = [vp]
    lib.b2d_exec_end_capture.argtypes = [vp, i]
    lib.b2d_exec_graph_nodes.argtypes = [vp, i]
    lib.b2d_exec_graph_nodes.restype = C.c_uint64
    lib.b2d_exec_run.argtypes = [vp, vp, sz, vp, sz, i64, i64, i64, i, C.POINTER(C.c_float)]
    lib.b2d_exec_sync.argtypes = [vp]
    lib.b2d_exec_steps_done.argtypes = [vp]
    lib.b2d_exec_steps_done.restype = C.c_uint64
    lib._exec_declared = True


class Trainer:
    def __init__(self, model, strategy, batch_per_replica: int, input_shape, use_graph: bool = True):
        self.model = model
        self.strategy = strategy
        self.batch = int(batch_per_replica)
        self.input_shape = tuple(input_shape)          # (H, W, C)
        self.layers = model.layers
        self.layout = ArenaLayout(self.layers)
        self.lr = float(model.optimizer.learning_rate)
        self.backend_name = strategy.backend_name
        if self.backend_name == "cuda":
            from .cuda_backend import CudaBackend
            from ..keras import mixed_precision
            torch.cuda.set_device(strategy.device)
            self.policy = mixed_precision.global_policy()
            self.be = CudaBackend(strategy.device, torch.bfloat16 if self.policy.compute_dtype == "bfloat16"
                                  else torch.float32)
        else:
            from .cpu_backend import CpuBackend
            self.be = CpuBackend()
            self._cpu_thread_layout(strategy)
        self.comm = strategy.make_comm(self.layout.n_total, self.be)
        self.params = self.comm.alloc(self.layout.n_total, self.be)
        self.use_graph = bool(use_graph and self.backend_name == "cuda" and self.comm.capturable)
        # Gradient arena(s): communicators that ping-pong gradients get one per graph slot.
        n_g = self.comm.num_grad_buffers if self.use_graph else 1
        self.grad_bufs = [self.comm.alloc(self.layout.n_total, self.be) for _ in range(n_g)]
        self.grads = self.grad_bufs[0]
        init = torch.from_numpy(self.layout.initial_values(self.layers, strategy.seed))
        self.params.copy_(init)
        for g in self.grad_bufs:
            g.zero_()
        if self.comm.world > 1 and not self.comm.all_equal(self.params):
            raise RuntimeError("replicas were not initialised identically (mirrored variables differ)")
        self._exec = None
        self._build()

    @staticmethod
    def _cpu_thread_layout(strategy):
        """Co-located CPU workers (the reference's multi-worker run on one host, BASELINE config 1) share the
        cores: worker r gets cores // world pool threads pinned to its own slice, instead of every worker
        spawning one thread per core.  B200DIST_CPU_THREADS overrides the per-worker thread count."""
        import os
        from .._native import cpu_lib
        world, rank = int(strategy.num_replicas_in_sync), int(getattr(strategy, "rank", 0))
        cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
        threads = int(os.environ.get("B200DIST_CPU_THREADS", "0")) or max(1, min(8, cores // max(world, 1)))
        if world > 1 or "B200DIST_CPU_THREADS" in os.environ:
            # slot 0 of the slice stays free for the (unpinned) calling thread: workers start one further on
            cpu_lib.lib().b2c_set_thread_layout(threads, (rank * threads) % max(cores, 1))

    # ------------------------------------------------------------------------------
    def _io_buffers(self):
        x = self.be.alloc((self.batch, *self.input_shape))
        y = self.be.alloc((self.batch,), torch.int32)
        return x, y

    def _build(self):
        lay, be = self.layout, self.be
        self.x_in, self.y_in, self.progs, self.updates = [], [], [], []
        n_slots = 2 if self.use_graph else 1
        for _ in range(n_slots):
            x, y = self._io_buffers()
            self.x_in.append(x)
            self.y_in.append(y)
        self.x_bytes = self.x_in[0].numel() * 4
        self.y_bytes = self.y_in[0].numel() * 4
        if self.backend_name != "cuda":
            self.progs.append(lower(self.layers, self.batch, be, lay, self.params, self.grads,
                                    self.x_in[0], self.y_in[0], lr=self.lr))
            self.updates.append(self.comm.make_update(be, self.params, self.grads, lay.n_params, self.lr))
            return
        from .._native import cuda_lib as nl
        lib = nl.lib()
        _declare_exec(lib)
        self._lib = lib
        h = C.c_void_p()
        nl.check(lib.b2d_exec_create(C.byref(h)), "b2d_exec_create")
        self._exec = h
        self.stream = torch.cuda.ExternalStream(lib.b2d_exec_stream(h), device=self.be.device)
        # metrics ring: pinned host memory the executor fills after every step
        self.ring_len = 4096
        self.metrics_host = torch.zeros((self.ring_len, TAIL), dtype=torch.float32).pin_memory()
        tail = self.params[lay.n_params:lay.n_total]
        nl.check(lib.b2d_exec_set_metrics(h, tail.data_ptr(), TAIL * 4, self.metrics_host.data_ptr(),
                                          self.ring_len), "b2d_exec_set_metrics")
        self.side_stream = torch.cuda.ExternalStream(lib.b2d_exec_side_stream(h), device=self.be.device)
        self.branch_stream = torch.cuda.Stream(device=self.be.device)     # shortcut branches of residual blocks
        self._capture_events = []                                        # fork / join events of the captures
        self.pipelined = []          # per slot: {"small", "deferred", "flush"} or empty
        self.bucketed = []           # per slot: {"early", "late"} or empty (intra-step overlap, default)
        torch.cuda.synchronize(self.be.device)
        import os
        with torch.cuda.stream(self.stream):
            for slot in range(n_slots):
                g_slot = self.grad_bufs[slot % len(self.grad_bufs)]
                prog = lower(self.layers, self.batch, be, lay, self.params, g_slot,
                             self.x_in[slot], self.y_in[slot], lr=self.lr)
                if len(self.grad_bufs) > 1:
                    g_other = self.grad_bufs[(slot + 1) % 2]
                    upd = self.comm.make_update(be, self.params, g_slot, lay.n_params, self.lr,
                                                grads_other=g_other)
                    # Optional (B200DIST_PIPELINE=1) bucketed, cross-step-overlapped update: everything
                    # but the first layer's parameters is reduced on a side branch while the NEXT
                    # step's first kernels run.  Measured SLOWER for the README model (the overlap
                    # window — one 5 us kernel — is shorter than the bucket's barrier + transfer), so
                    # it is off by default; it pays once the first layer's forward is long.
                    # Optional (B200DIST_BUCKETS=1): two buckets inside the step.  Everything after the first
                    # parametrised layer is final once that layer's own backward starts, so its fused
                    # all-reduce+SGD runs on the side branch UNDER the first layer's backward; only the first
                    # layer's (small) bucket and the metric sums remain on the critical path.  Measured on the
                    # README model: 57.1 vs 53.6 us/step at 2 GPUs, 66.2 vs 65.3 us at 8 — the second kernel's
                    # cross-GPU barrier costs what the overlap hides, so one fused update per step stays the default.
                    bucket_begin = self._bucket_begin(prog)
                    if (self.use_graph and bucket_begin is not None and hasattr(self.comm, "make_bucketed_updates")
                            and os.environ.get("B200DIST_BUCKETS", "0") == "1"
                            and os.environ.get("B200DIST_PIPELINE", "0") != "1"):
                        self.bucketed.append(self.comm.make_bucketed_updates(
                            be, self.params, g_slot, g_other, lay.n_params, self.lr, bucket_begin))
                    deferred_begin = self._deferred_begin(prog)
                    if (self.use_graph and deferred_begin is not None
                            and hasattr(self.comm, "make_pipelined_updates")
                            and os.environ.get("B200DIST_PIPELINE", "0") == "1"):
                        self.pipelined.append(self.comm.make_pipelined_updates(
                            be, self.params, g_slot, g_other, lay.n_params, self.lr, deferred_begin))
                else:
                    upd = self.comm.make_update(be, self.params, g_slot, lay.n_params, self.lr)
                self.progs.append(prog)
                self.updates.append(upd)
                nl.check(lib.b2d_exec_set_io(h, slot, self.x_in[slot].data_ptr(), self.x_bytes,
                                             self.y_in[slot].data_ptr(), self.y_bytes), "b2d_exec_set_io")
            if self.use_graph:
                # warm every kernel once outside capture (module load, smem attributes), on zero
                # inputs with the update skipped so the weights are untouched
                self.progs[0].run()
                self.grad_bufs[0].zero_()
                self.stream.synchronize()
                for slot in range(n_slots):
                    prog = self.progs[slot]
                    nl.check(lib.b2d_exec_begin_capture(h), "b2d_exec_begin_capture")
                    if len(self.pipelined) == n_slots:
                        pipe = self.pipelined[slot]
                        nl.check(lib.b2d_exec_fork(h), "b2d_exec_fork")
                        with torch.cuda.stream(self.side_stream):
                            pipe["deferred"]()            # previous step's big bucket
                        for f in prog.forward[:prog.early_forward]:
                            f()                            # needs only the first layer's weights
                        nl.check(lib.b2d_exec_join(h), "b2d_exec_join")
                        for f in prog.forward[prog.early_forward:]:
                            f()
                        for f in prog.backward:
                            f()
                        pipe["small"]()                    # the only communication on the critical path
                    elif len(self.bucketed) == n_slots:
                        self._run_step_branches(prog, early=self.bucketed[slot]["early"])
                        self.bucketed[slot]["late"]()
                    else:
                        self._run_step_branches(prog)
                        self.updates[slot]()
                    nl.check(lib.b2d_exec_end_capture(h, slot), "b2d_exec_end_capture")
                self.kernels_per_step = int(lib.b2d_exec_graph_nodes(h, 0))
            else:
                self.kernels_per_step = None
        self.stream.synchronize()

    def _run_step_branches(self, prog, early=None):
        """Issue one step for capture.  Weight-gradient launches (tagged ``side_ok`` by the lowering) go to
        the executor's side stream, forked right where they sit in the backward list and joined before the
        update: in the graph they become a branch that runs concurrently with the data-gradient chain
        (B200DIST_SIDE_WGRAD=0 keeps the step a single chain).  Launches tagged ``branch`` (the shortcut of a
        residual block, forward and backward) run on a third stream from their first launch to their first
        consumer (``join_branch``); B200DIST_SIDE_SHORTCUT=0 keeps them on the main stream.  ``early``: the fused all-reduce+SGD of every
        layer after the first parametrised one; it is queued on the side stream — behind the weight gradients
        that feed it and after the main stream's data gradients that still read those weights — at the point
        where only the first layer's backward remains."""
        import os
        from .._native import cuda_lib as nl
        side = os.environ.get("B200DIST_SIDE_WGRAD", "1") != "0"
        use_branch = os.environ.get("B200DIST_SIDE_SHORTCUT", "1") != "0"
        main = torch.cuda.current_stream()
        state = {"open": False}

        def join_branch():
            if state["open"]:
                ev = torch.cuda.Event()
                ev.record(self.branch_stream)
                main.wait_event(ev)
                self._capture_events.append(ev)
                state["open"] = False

        def on_branch(f):
            """Shortcut branch of a residual block: its own stream, forked where its first launch sits."""
            if not state["open"]:
                ev = torch.cuda.Event()
                ev.record(main)
                self.branch_stream.wait_event(ev)
                self._capture_events.append(ev)
                state["open"] = True
            with torch.cuda.stream(self.branch_stream):
                f()

        for f in prog.forward:
            if getattr(f, "join_branch", False):
                join_branch()
            if use_branch and getattr(f, "branch", False):
                on_branch(f)
            else:
                f()
        join_branch()
        forked = False
        for i, f in enumerate(prog.backward):
            if getattr(f, "join_branch", False):
                join_branch()
            if early is not None and i == prog.early_backward:
                nl.check(self._lib.b2d_exec_fork(self._exec), "b2d_exec_fork")
                with torch.cuda.stream(self.side_stream):
                    early()
                forked, side = True, False      # later weight gradients stay on the main stream
            if use_branch and getattr(f, "branch", False):
                on_branch(f)                    # incl. the branch's own weight gradients: they read its dy
            elif side and getattr(f, "side_ok", False):
                nl.check(self._lib.b2d_exec_fork(self._exec), "b2d_exec_fork")
                with torch.cuda.stream(self.side_stream):
                    f()
                forked = True
            else:
                f()
        join_branch()
        if forked:
            nl.check(self._lib.b2d_exec_join(self._exec), "b2d_exec_join")

    def _bucket_begin(self, prog):
        """Arena offset where the early bucket starts (end of the first parametrised layer), or None."""
        if prog.first_param_layer < 0 or not 0 <= prog.early_backward < len(prog.backward):
            return None
        later = [s for s in self.layout.slots if s.layer_index > prog.first_param_layer]
        if not later:
            return None
        begin = min(s.offset for s in later)
        return begin if 0 < begin < self.layout.n_params else None

    def _deferred_begin(self, prog):
        """Arena offset where the deferred bucket starts (end of the first parametrised layer), or
        None when the model has nothing to defer."""
        if prog.first_param_layer < 0 or prog.early_forward <= 0:
            return None
        later = [s for s in self.layout.slots if s.layer_index > prog.first_param_layer]
        if not later:
            return None
        begin = min(s.offset for s in later)
        return begin if 0 < begin < self.layout.n_params else None

    def _flush_pending(self):
        """Apply the deferred bucket of the last executed step (all replicas call this together)."""
        if self.pipelined and self._pending_slot is not None:
            with torch.cuda.stream(self.stream):
                self.pipelined[self._pending_slot]["flush"]()
            self._pending_slot = None

    _pending_slot = None

    # ------------------------------------------------------------------------------
    def describe(self) -> list[str]:
        return list(self.progs[0].description)

    def run_steps(self, x_src, y_src, first_step: int, steps_in_src: int, n_steps: int,
                  global_batch: int, timed: bool = False):
        """Run ``n_steps`` steps.  ``x_src`` [S,H,W,C] fp32 and ``y_src`` [S] int32 are pinned
        host or device tensors; step t of this replica reads rows
        ``[(t % steps_in_src) * global_batch + rank * B, ... + B)``.  Returns elapsed ms if timed."""
        B, r = self.batch, self.comm.rank
        row = int(np.prod(self.input_shape))
        if self._exec is not None and self.use_graph:
            xs = x_src.data_ptr() + r * B * row * 4
            ys = y_src.data_ptr() + r * B * 4
            from .._native import cuda_lib as nl
            if timed:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(self.stream)
            nl.check(self._lib.b2d_exec_run(self._exec, xs, global_batch * row * 4, ys, global_batch * 4,
                                            first_step, steps_in_src, n_steps, 0, None), "b2d_exec_run")
            if n_steps > 0 and self.pipelined:
                self._pending_slot = (self.steps_done() - 1) & 1
                self._flush_pending()          # inside the timed region: the work is not skipped
            if timed:
                e1.record(self.stream)
                e1.synchronize()
                return e0.elapsed_time(e1)
            self.sync()
            return None
        # eager loops (CPU backend; CUDA with a non-capturable communicator)
        t0 = time.perf_counter()
        cuda = self.backend_name == "cuda"
        ctx = torch.cuda.stream(self.stream) if cuda else _null()
        if cuda and timed:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with ctx:
            if cuda and timed:
                e0.record()
            for t in range(n_steps):
                s = (first_step + t) % steps_in_src
                lo = s * global_batch + r * B
                if cuda:
                    self.x_in[0].copy_(x_src[lo:lo + B], non_blocking=True)
                    self.y_in[0].copy_(y_src[lo:lo + B], non_blocking=True)
                else:
                    # plain memmove: torch's copy_ wakes its OpenMP team, whose spin-wait then competes with
                    # the C++ backend's own worker pool for the same cores (conv forward 0.7 -> 4 ms)
                    C.memmove(self.x_in[0].data_ptr(), x_src.data_ptr() + lo * row * 4, B * row * 4)
                    C.memmove(self.y_in[0].data_ptr(), y_src.data_ptr() + lo * 4, B * 4)
                self.progs[0].run()
                self.updates[0]()
            if cuda and timed:
                e1.record()
                e1.synchronize()
                return e0.elapsed_time(e1)
            if cuda:
                self.stream.synchronize()
        return (time.perf_counter() - t0) * 1e3 if timed else None

    def sync(self):
        if self._exec is not None:
            from .._native import cuda_lib as nl
            nl.check(self._lib.b2d_exec_sync(self._exec), "b2d_exec_sync")

    def steps_done(self) -> int:
        return int(self._lib.b2d_exec_steps_done(self._exec)) if self._exec is not None else 0

    def metric_sums(self) -> tuple[float, float, float]:
        """Running (loss_sum, correct, samples) since the last reset, summed over replicas."""
        self.sync()
        t = self.params[self.layout.n_params:self.layout.n_total].detach().cpu()
        return float(t[0]), float(t[1]), float(t[2])

    def step_metrics(self, n_last: int) -> np.ndarray:
        """The last ``n_last`` per-step snapshots of the running sums (CUDA graph mode)."""
        if self._exec is None or not self.use_graph:
            return np.zeros((0, TAIL), np.float32)
        self.sync()
        done = self.steps_done()
        n = min(n_last, done, self.ring_len)
        idx = [(done - n + k) % self.ring_len for k in range(n)]
        return self.metrics_host.numpy()[idx].copy()

    def reset_metrics(self):
        self.sync()
        self.params[self.layout.n_params:self.layout.n_total].zero_()
        if self.backend_name == "cuda":
            torch.cuda.synchronize(self.be.device)
        self.comm.barrier()

    def get_weights(self) -> list[np.ndarray]:
        self.sync()
        flat = self.params.detach().cpu()
        return [self.layout.view(flat, s).numpy().copy() for s in self.layout.slots]

    def set_weights(self, arrays):
        self.sync()
        assert len(arrays) == len(self.layout.slots), "wrong number of weight arrays"
        flat = self.params.detach().cpu().clone()
        for s, a in zip(self.layout.slots, arrays):
            a = np.asarray(a, np.float32)
            if tuple(a.shape) != s.shape:
                raise ValueError(f"{s.name}: expected shape {s.shape}, got {a.shape}")
            flat[s.offset:s.offset + s.size] = torch.from_numpy(a.reshape(-1))
        self.params.copy_(flat)
        if self.backend_name == "cuda":
            torch.cuda.synchronize(self.be.device)

    def forward_logits(self, x: torch.Tensor) -> torch.Tensor:
        """Inference on one batch of exactly ``batch`` rows (used by predict / evaluate)."""
        if not hasattr(self, "_infer"):
            xi, yi = self._io_buffers()
            ctx = torch.cuda.stream(self.stream) if self.backend_name == "cuda" else _null()
            with ctx:
                self._infer = (xi, lower(self.layers, self.batch, self.be, self.layout, self.params,
                                         self.grads, xi, yi, training=False))
        xi, prog = self._infer
        ctx = torch.cuda.stream(self.stream) if self.backend_name == "cuda" else _null()
        with ctx:
            xi.copy_(x)
            for f in prog.forward:
                f()
            out = prog.logits.detach().clone()
        if self.backend_name == "cuda":
            self.stream.synchronize()
        return out.cpu()

    def close(self):
        if self._exec is not None:
            self.sync()
            self._lib.b2d_exec_destroy(self._exec)
            self._exec = None
        self.comm.close()


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False
