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
    lib.b2d_exec_set_metrics_slot.argtypes = [vp, i, vp]
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
        self._exec = None
        self.comm = strategy.make_comm(self.layout.n_total, self.be)
        self.use_graph = bool(use_graph and self.backend_name == "cuda" and self.comm.capturable)
        # Arenas.  Communicators that overlap the all-reduce with the backward pass (SymmComm) ping-pong both the
        # gradients and the weights between two buffers, one per graph slot: step t reads W[t & 1], accumulates
        # into G[t & 1], and its update writes W[(t + 1) & 1].  Everyone else has one buffer of each.
        if hasattr(self.comm, "alloc_arenas"):
            self.param_bufs, self.grad_bufs = self.comm.alloc_arenas(self.layout.n_total, self.be, self.use_graph)
        else:
            self.param_bufs = [self.comm.alloc(self.layout.n_total, self.be)]
            n_g = self.comm.num_grad_buffers if self.use_graph else 1
            self.grad_bufs = [self.comm.alloc(self.layout.n_total, self.be) for _ in range(n_g)]
        self.grads = self.grad_bufs[0]
        init = torch.from_numpy(self.layout.initial_values(self.layers, strategy.seed))
        for w in self.param_bufs:
            w.copy_(init)
        for g in self.grad_bufs:
            g.zero_()
        if self.comm.world > 1 and not self.comm.all_equal(self.params):
            raise RuntimeError("replicas were not initialised identically (mirrored variables differ)")
        self._exec = None
        self._build()

    @property
    def params(self):
        """The weight arena holding the CURRENT weights (with weight ping-pong: the buffer the next step reads)."""
        if len(self.param_bufs) == 1:
            return self.param_bufs[0]
        return self.param_bufs[self.steps_done() & 1]

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
        tail = self.param_bufs[0][lay.n_params:lay.n_total]
        nl.check(lib.b2d_exec_set_metrics(h, tail.data_ptr(), TAIL * 4, self.metrics_host.data_ptr(),
                                          self.ring_len), "b2d_exec_set_metrics")
        if len(self.param_bufs) == 2:
            # the step replayed from slot s leaves the running sums in the tail of W[1 - s]
            for slot in range(2):
                t = self.param_bufs[1 - slot][lay.n_params:lay.n_total]
                nl.check(lib.b2d_exec_set_metrics_slot(h, slot, t.data_ptr()), "b2d_exec_set_metrics_slot")
        self.side_stream = torch.cuda.ExternalStream(lib.b2d_exec_side_stream(h), device=self.be.device)
        self.branch_stream = torch.cuda.Stream(device=self.be.device)     # shortcut branches of residual blocks
        self._capture_events = []                                        # fork / join events of the captures
        self.schedules = []          # per slot: (list of (n_backward_launches, launch), closing launch) or None
        self.comm_stream = torch.cuda.Stream(device=self.be.device)      # overlapped gradient buckets
        torch.cuda.synchronize(self.be.device)
        with torch.cuda.stream(self.stream):
            for slot in range(n_slots):
                g_slot = self.grad_bufs[slot % len(self.grad_bufs)]
                w_slot = self.param_bufs[slot % len(self.param_bufs)]
                overlapped = len(self.param_bufs) == 2
                # GEMM -> reduce-scatter: with the overlapped update, large Dense weight-gradient GEMMs add their
                # result straight into the owning replica's gradient arena over NVLink (cuda_backend.dense_wgrad)
                be.scattered = []
                be.scatter_target = self.comm.scatter_target(g_slot) if overlapped else None
                prog = lower(self.layers, self.batch, be, lay, w_slot, g_slot,
                             self.x_in[slot], self.y_in[slot], lr=self.lr)
                prog.scattered = sorted(be.scattered)
                be.scatter_target = None
                sched = None
                if len(self.param_bufs) == 2:
                    # Default for peer-memory replicas: the update is a bucket schedule overlapped with the
                    # backward pass (see _bucket_plan); B200DIST_BUCKETS=0 falls back to one fused launch.
                    g_other, w_other = self.grad_bufs[1 - slot], self.param_bufs[1 - slot]
                    early, late_segs = self._bucket_plan(prog)
                    fns, late, desc = self.comm.make_step_updates(
                        be, w_slot, w_other, g_slot, g_other, lay.n_params, self.lr,
                        [(b, e, sc) for _, b, e, sc in early], late_segs)
                    sched = ([(n, f) for (n, _, _, _), f in zip(early, fns)], late)
                    upd = late
                    self.update_schedule = desc
                elif len(self.grad_bufs) > 1:
                    g_other = self.grad_bufs[(slot + 1) % 2]
                    upd = self.comm.make_update(be, w_slot, g_slot, lay.n_params, self.lr, grads_other=g_other)
                else:
                    upd = self.comm.make_update(be, w_slot, g_slot, lay.n_params, self.lr)
                self.schedules.append(sched)
                self.progs.append(prog)
                self.updates.append(upd)
                nl.check(lib.b2d_exec_set_io(h, slot, self.x_in[slot].data_ptr(), self.x_bytes,
                                             self.y_in[slot].data_ptr(), self.y_bytes), "b2d_exec_set_io")
            if self.use_graph:
                # warm every kernel once outside capture (module load, smem attributes), on zero
                # inputs with the update skipped so the weights are untouched
                self.progs[0].run()
                self.stream.synchronize()
                if getattr(self.progs[0], "scattered", None):
                    self.comm.barrier()           # every replica's warm-up pushes into the peers' arenas have landed
                self.grad_bufs[0].zero_()
                self.stream.synchronize()
                if getattr(self.progs[0], "scattered", None):
                    self.comm.barrier()           # nobody starts before every arena is clean again
                for slot in range(n_slots):
                    prog = self.progs[slot]
                    nl.check(lib.b2d_exec_begin_capture(h), "b2d_exec_begin_capture")
                    self._run_step_branches(prog, self.schedules[slot])
                    self.updates[slot]()
                    nl.check(lib.b2d_exec_end_capture(h, slot), "b2d_exec_end_capture")
                self.kernels_per_step = int(lib.b2d_exec_graph_nodes(h, 0))
            else:
                self.kernels_per_step = None
        self.stream.synchronize()
        if self.comm.world > 1:
            # graph capture / instantiation takes a different time on every replica; meet here so that the first
            # step's device-side waits start together (they are bounded by B200DIST_AR_TIMEOUT_S)
            self.comm.barrier()

    def _run_step_branches(self, prog, sched=None):
        """Issue one step for capture.  Weight-gradient launches (tagged ``side_ok`` by the lowering) go to
        the executor's side stream, forked right where they sit in the backward list and joined before the
        update: in the graph they become a branch that runs concurrently with the data-gradient chain
        (B200DIST_SIDE_WGRAD=0 keeps the step a single chain).  Launches tagged ``branch`` (the shortcut of a
        residual block, forward and backward) run on a third stream from their first launch to their first
        consumer (``join_branch``); B200DIST_SIDE_SHORTCUT=0 keeps them on the main stream.

        ``sched`` = (early buckets, closing launch): every early bucket ``(n, launch)`` — a fused
        all-reduce+SGD over gradients that are final once the first ``n`` backward launches are done — is
        queued on the communication stream at that point, behind everything issued so far on the main, side and
        shortcut streams, and runs under the rest of the backward pass.  The weights it writes belong to the
        OTHER weight buffer, so nothing in this step reads them.  All branches join before the closing launch."""
        import os
        from .._native import cuda_lib as nl
        side = os.environ.get("B200DIST_SIDE_WGRAD", "1") != "0"
        use_branch = os.environ.get("B200DIST_SIDE_SHORTCUT", "1") != "0"
        main = torch.cuda.current_stream()
        state = {"open": False, "forked": False, "comm": False}
        early = list(sched[0]) if sched else []

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

        # index of the last weight-gradient launch that goes to the side stream
        last_side = max((i for i, f in enumerate(prog.backward) if side and getattr(f, "side_ok", False)
                         and not (use_branch and getattr(f, "branch", False))), default=-1)
        on_side = os.environ.get("B200DIST_BUCKET_ON_SIDE", "1") != "0"

        def issue_buckets(n_done):
            while early and early[0][0] <= n_done:
                _, launch = early.pop(0)
                if (on_side and state["forked"] and not state["open"] and not state["comm"]
                        and n_done - 1 == last_side and not early):
                    # The bucket became final with the LAST weight-gradient launch of the step and nothing else
                    # will use the side stream: queue it right behind that launch on the same stream.  The edge
                    # is then a programmatic one (the kernel is resident and waiting when its producer ends)
                    # instead of a cross-stream dependency, which costs a launch latency on the critical path.
                    with torch.cuda.stream(self.side_stream):
                        launch()
                    continue
                streams = [main] + ([self.side_stream] if state["forked"] else []) + \
                          ([self.branch_stream] if state["open"] else [])
                for st in streams:
                    ev = torch.cuda.Event()
                    ev.record(st)
                    self.comm_stream.wait_event(ev)
                    self._capture_events.append(ev)
                with torch.cuda.stream(self.comm_stream):
                    launch()
                state["comm"] = True

        for f in prog.forward:
            if getattr(f, "join_branch", False):
                join_branch()
            if use_branch and getattr(f, "branch", False):
                on_branch(f)
            else:
                f()
        join_branch()
        issue_buckets(0)
        for i, f in enumerate(prog.backward):
            if getattr(f, "join_branch", False):
                join_branch()
            if use_branch and getattr(f, "branch", False):
                on_branch(f)                    # incl. the branch's own weight gradients: they read its dy
            elif side and getattr(f, "side_ok", False):
                nl.check(self._lib.b2d_exec_fork(self._exec), "b2d_exec_fork")
                with torch.cuda.stream(self.side_stream):
                    f()
                state["forked"] = True
            else:
                f()
            issue_buckets(i + 1)
        join_branch()
        if state["forked"]:
            nl.check(self._lib.b2d_exec_join(self._exec), "b2d_exec_join")
        if state["comm"]:
            ev = torch.cuda.Event()
            ev.record(self.comm_stream)
            main.wait_event(ev)
            self._capture_events.append(ev)

    def _bucket_plan(self, prog):
        """Cut the gradient arena into buckets for the overlapped update.

        Returns (early, late_segments).  ``early`` = [(n, begin, end, scattered)]: the floats [begin, end) are final
        once the first n backward launches are done (``scattered``: already summed into their owners' arenas by the
        producing GEMM); buckets run from the last layer towards the first, each at least
        B200DIST_BUCKET_MB (default 16 MB) unless the model is smaller, whole top-level layers only.
        ``late_segments``: what only becomes final when the backward pass ends — the first parametrised layer —
        plus the metric tail; it is the closing launch of the step."""
        import os
        lay = self.layout
        n_bwd = len(prog.backward)
        target = max(1, int(float(os.environ.get("B200DIST_BUCKET_MB", "16")) * (1 << 20) / 4))
        table = list(prog.grad_final)
        cuts, end = [], lay.n_params
        rest = None                         # (n, begin) of the lowest layer that may still join an early bucket
        for begin, n in reversed(table[1:]):
            if n >= n_bwd or begin <= 0:
                break
            rest = (n, begin)
            if end - begin >= target:
                cuts.append((n, begin, end))
                end, rest = begin, None
        if rest is not None and rest[1] < end:
            cuts.append((rest[0], rest[1], end))
            end = rest[1]
        # Ranges that their weight-gradient GEMM has already reduce-scattered become buckets of their own
        # (the update only has to finish them); what is left of a bucket around them is reduced as before.
        early = []
        for n, b, e in cuts:
            pos = b
            for sb, se in getattr(prog, "scattered", []):
                if sb < b or se > e:
                    continue
                if sb > pos:
                    early.append((n, pos, sb, False))
                early.append((n, sb, se, True))
                pos = se
            if pos < e:
                early.append((n, pos, e, False))
        early.sort(key=lambda t: t[0])
        late = [(0, end), (lay.n_params, lay.n_total)]
        # A small plain remainder that touches the metric tail (biases and the last layers of a small model)
        # rides in the closing exchange instead of paying for a launch and a barrier of its own.
        ll_max = getattr(self.comm, "ll_max_floats", 0)
        tails = [t for t in early if not t[3] and t[2] == lay.n_params]
        if tails and (end + lay.n_total - tails[0][1]) <= ll_max:
            early.remove(tails[0])
            late = [(0, end), (tails[0][1], lay.n_total)]
        return early, late

    # ------------------------------------------------------------------------------
    def describe(self) -> list[str]:
        return list(self.progs[0].description)

    def run_steps(self, x_src, y_src, first_step: int, steps_in_src: int, n_steps: int,
                  global_batch: int, timed: bool = False, pre_steps: int = 0):
        """Run ``n_steps`` steps.  ``x_src`` [S,H,W,C] fp32 and ``y_src`` [S] int32 are pinned
        host or device tensors; step t of this replica reads rows
        ``[(t % steps_in_src) * global_batch + rank * B, ... + B)``.  Returns elapsed ms if timed.

        ``pre_steps`` (timed runs): that many extra steps are enqueued on the same stream BEFORE the start
        event, with no host synchronisation in between.  The fused all-reduce of those steps lines the
        replicas up on the device, so the timed region starts with every GPU in lock step instead of
        absorbing the host-side release skew of the processes (milliseconds at 8 ranks)."""
        B, r = self.batch, self.comm.rank
        row = int(np.prod(self.input_shape))
        if self._exec is not None and self.use_graph:
            xs = x_src.data_ptr() + r * B * row * 4
            ys = y_src.data_ptr() + r * B * 4
            from .._native import cuda_lib as nl
            if timed:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                if pre_steps > 0:
                    nl.check(self._lib.b2d_exec_run(self._exec, xs, global_batch * row * 4, ys, global_batch * 4,
                                                    first_step, steps_in_src, pre_steps, 0, None), "b2d_exec_run")
                    first_step += pre_steps
                e0.record(self.stream)
            nl.check(self._lib.b2d_exec_run(self._exec, xs, global_batch * row * 4, ys, global_batch * 4,
                                            first_step, steps_in_src, n_steps, 0, None), "b2d_exec_run")
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
        if self._exec is not None and self.use_graph and self.steps_done() > getattr(self, "_reset_at", -1) >= 0:
            # the executor streamed the sums of every step into the pinned ring: no extra device read
            t = self.metrics_host[(self.steps_done() - 1) % self.ring_len]
            return float(t[0]), float(t[1]), float(t[2])
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
        """Zero the running sums (start of an epoch).  The sums live in the tail of this replica's own weight
        arena(s) and only this replica's kernels write them, so no cross-replica synchronisation is involved; in
        graph mode the clear is queued on the executor's stream, ordered before the next step, with no host wait."""
        if self._exec is not None and self.use_graph:
            with torch.cuda.stream(self.stream):
                for w in self.param_bufs:
                    w[self.layout.n_params:self.layout.n_total].zero_()
            self._reset_at = self.steps_done()
            return
        self.sync()
        for w in self.param_bufs:
            w[self.layout.n_params:self.layout.n_total].zero_()
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
        for w in self.param_bufs:
            w.copy_(flat)
        if self.backend_name == "cuda":
            torch.cuda.synchronize(self.be.device)

    def forward_logits(self, x: torch.Tensor) -> torch.Tensor:
        """Inference on one batch of exactly ``batch`` rows (used by predict / evaluate)."""
        if not hasattr(self, "_infer"):
            self._infer = {}
        self.sync()
        cur = (self.steps_done() & 1) if len(self.param_bufs) == 2 else 0      # the buffer with the current weights
        if cur not in self._infer:
            xi, yi = self._io_buffers()
            ctx = torch.cuda.stream(self.stream) if self.backend_name == "cuda" else _null()
            with ctx:
                self._infer[cur] = (xi, lower(self.layers, self.batch, self.be, self.layout, self.param_bufs[cur],
                                              self.grads, xi, yi, training=False))
        xi, prog = self._infer[cur]
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
