"""SymmComm: the GPU communicator — fused all-reduce + average + SGD over NVLink peer memory.

Each replica owns one block of *symmetric memory* (csrc/runtime/symm.cpp): two weight arenas, two gradient
arenas (both ping-pong with the two step graphs), signal pads and the LL inboxes at the same offsets on every GPU,
with every peer's block mapped into every process and, where the fabric exposes it, an NVLS multicast view.
The update of a step is a schedule of launches from csrc/cuda/allreduce.cu that read the peers' gradients and
write the peers' weights directly through NVSwitch: overlapped buckets (one-shot, two-shot, NVLS by size, or
"finish what the weight-gradient GEMM already reduce-scattered") plus a closing push exchange for the small last
bucket — all captured into the step's CUDA graph (engine/trainer.py decides what runs when).  No NCCL call
exists on this path; torch.distributed (gloo) is used only as the control plane (barriers around set-up,
replica-consistency checks).

This replaces the reference's "Collective batch_all_reduce: 6 all-reduces" + 4 metric
all-reduces + optimizer step per iteration (README.md:403-407, 302).
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from .._native import cuda_lib as nl

MAX_RANKS = 8
ALGOS = {"oneshot": 0, "one_shot": 0, "twoshot": 1, "two_shot": 1, "nvls": 2, "pull": 3, "auto": -1}
ALGO_NAMES = {0: "one-shot", 1: "two-shot", 2: "NVLS multicast", 3: "two-shot pull"}

_VP8 = C.c_void_p * MAX_RANKS


class AllreduceDesc(C.Structure):
    _fields_ = [("w", _VP8), ("g", _VP8), ("sig", _VP8), ("epoch", _VP8), ("out", _VP8),
                ("w_mc", C.c_void_p), ("g_mc", C.c_void_p),
                ("n_params", C.c_int64), ("n_total", C.c_int64), ("lr", C.c_float),
                ("rank", C.c_int), ("world", C.c_int), ("algo", C.c_int), ("emulate", C.c_int),
                ("plain", C.c_int), ("blocks", C.c_int), ("g_zero", C.c_void_p), ("timeline", C.c_void_p),
                ("seg_begin", C.c_int64 * 2), ("seg_end", C.c_int64 * 2), ("nseg", C.c_int),
                ("w_out", _VP8), ("w_out_mc", C.c_void_p), ("flags", C.c_int)]


class LlDesc(C.Structure):
    _fields_ = [("inbox", _VP8), ("epoch", _VP8), ("w", _VP8), ("w_out", _VP8), ("g", _VP8), ("g_zero", _VP8),
                ("seg_begin", C.c_int64 * 2), ("seg_end", C.c_int64 * 2), ("nseg", C.c_int),
                ("n_params", C.c_int64), ("lr", C.c_float), ("rank", C.c_int), ("world", C.c_int),
                ("emulate", C.c_int), ("timeline", C.c_void_p)]


AR_NO_EXIT_BARRIER, AR_STRICT_ENTRY = 1, 2


def _declare(lib):
    if getattr(lib, "_symm_declared", False):
        return
    vp, i, sz = C.c_void_p, C.c_int, C.c_size_t
    lib.b2d_symm_create.argtypes = [C.POINTER(vp), C.c_char_p, i, i, i, sz, C.c_double]
    lib.b2d_symm_multicast_phase1.argtypes = [vp]
    lib.b2d_symm_multicast_phase2.argtypes = [vp]
    lib.b2d_symm_ptr.argtypes = [vp, i]
    lib.b2d_symm_ptr.restype = vp
    lib.b2d_symm_multicast_ptr.argtypes = [vp]
    lib.b2d_symm_multicast_ptr.restype = vp
    lib.b2d_symm_bytes.argtypes = [vp]
    lib.b2d_symm_bytes.restype = sz
    lib.b2d_symm_multicast_supported.argtypes = [vp]
    lib.b2d_symm_destroy.argtypes = [vp]
    lib.b2d_allreduce_signal_bytes.restype = sz
    lib.b2d_allreduce_epoch_bytes.restype = sz
    lib.b2d_allreduce_pick_algo.argtypes = [C.c_int64, i, i]
    lib.b2d_allreduce_sgd.argtypes = [C.POINTER(AllreduceDesc), vp]
    lib.b2d_ll_inbox_bytes.restype = sz
    lib.b2d_ll_epoch_bytes.restype = sz
    lib.b2d_ll_max_floats.restype = C.c_int64
    lib.b2d_ll_allreduce_sgd.argtypes = [C.POINTER(LlDesc), vp]
    lib._symm_declared = True


class _Raw:
    """Zero-copy view of raw device memory for torch.as_tensor."""

    def __init__(self, ptr: int, n: int, typestr: str = "<f4"):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False),
                                         "version": 3, "strides": None}


def _align(n: int, a: int = 256) -> int:
    return (n + a - 1) // a * a


class SymmComm:
    name = "symm"
    capturable = True

    def __init__(self, strategy, tag: str, capacity_floats: int, device, timeout_s: float = 120.0):
        import torch.distributed as dist
        lib = nl.lib()
        _declare(lib)
        self._lib, self._dist = lib, dist
        self.rank, self.world = strategy.rank, strategy.num_workers
        if self.world > MAX_RANKS:
            raise RuntimeError(f"SymmComm supports up to {MAX_RANKS} replicas (one NVSwitch box)")
        self.device = torch.device(device)
        self.n = _align(capacity_floats, 64)
        arena_bytes = self.n * 4
        self.sig_bytes = int(lib.b2d_allreduce_signal_bytes())
        self.epoch_bytes = int(lib.b2d_allreduce_epoch_bytes())
        self.ll_inbox_bytes = int(lib.b2d_ll_inbox_bytes())
        self.ll_epoch_bytes = int(lib.b2d_ll_epoch_bytes())
        self.ll_max_floats = int(lib.b2d_ll_max_floats())
        # layout of the symmetric block (same offsets on every replica)
        a = _align(arena_bytes, 4096)
        self.off_w = 0
        self.off_g = a
        self.off_g2 = 2 * a                                        # second gradient buffer (ping-pong)
        self.off_w2 = 3 * a                                        # second weight buffer (ping-pong)
        self.off_sig = 4 * a
        self.off_epoch = self.off_sig + _align(self.sig_bytes * self.CHANNELS, 4096)
        self.off_ll = self.off_epoch + _align(self.epoch_bytes * self.CHANNELS, 4096)
        self.off_ll_epoch = self.off_ll + _align(self.ll_inbox_bytes, 4096)
        total = self.off_ll_epoch + _align(self.ll_epoch_bytes, 4096)
        h = C.c_void_p()
        torch.cuda.set_device(self.device)
        dist.barrier()
        nl.check(lib.b2d_symm_create(C.byref(h), tag.encode(), self.rank, self.world, self.device.index,
                                     total, timeout_s), "b2d_symm_create")
        self._h = h
        self.base = [int(lib.b2d_symm_ptr(h, p)) for p in range(self.world)]
        dist.barrier()
        # NVLS multicast where the fabric exposes it (optional; agreed on by all replicas)
        self.mc_base = 0
        self.mc_reason = "disabled by B200DIST_NVLS=0"
        want = os.environ.get("B200DIST_NVLS", "1") != "0"
        if want:
            self.mc_reason = "CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED is 0"
        ok = torch.tensor([1 if (want and lib.b2d_symm_multicast_supported(h)) else 0])
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok[0]):
            rc = lib.b2d_symm_multicast_phase1(h)
            if rc != 0:
                self.mc_reason = "create/add-device failed: " + lib.b2d_last_error().decode(errors="replace")
            ok = torch.tensor([1 if rc == 0 else 0])
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if int(ok[0]):
                rc = lib.b2d_symm_multicast_phase2(h)
                if rc != 0:
                    self.mc_reason = "bind/map failed: " + lib.b2d_last_error().decode(errors="replace")
                ok = torch.tensor([1 if rc == 0 else 0])
                dist.all_reduce(ok, op=dist.ReduceOp.MIN)
                if int(ok[0]):
                    self.mc_base = int(lib.b2d_symm_multicast_ptr(h))
                    self.mc_reason = "bound"
            elif rc == 0:
                self.mc_reason = "a peer failed to set up multicast"
        self._next = 0
        mode = os.environ.get("B200DIST_ALLREDUCE", "").lower() or {
            "ONE_SHOT": "oneshot", "TWO_SHOT": "twoshot", "NVLS": "nvls"}.get(strategy.communication.name, "auto")
        if mode not in ALGOS:
            raise ValueError(f"B200DIST_ALLREDUCE must be one of {sorted(ALGOS)}")
        self.algo = ALGOS[mode]
        if self.algo == 2 and not self.mc_base:
            raise RuntimeError(f"NVLS all-reduce requested but multicast is not available on this box ({self.mc_reason})")
        resolved = self.algo if self.algo >= 0 else int(
            lib.b2d_allreduce_pick_algo(self.n, self.world, 1 if self.mc_base else 0))
        self.algo_resolved = resolved
        self.description = (f"fused peer-memory all-reduce+SGD over NVLink ({ALGO_NAMES[resolved]}, "
                            f"{self.world} GPUs, multicast {'bound' if self.mc_base else 'unavailable: ' + self.mc_reason})")

    # -- symmetric tensors -------------------------------------------------------------
    def _tensor(self, off_bytes: int, n: int, typestr="<f4"):
        t = torch.as_tensor(_Raw(self.base[self.rank] + off_bytes, n, typestr), device=self.device)
        t._b2d_owner = self       # the mapping lives as long as the communicator
        return t

    CHANNELS = 4              # independent signal pads: kernels of different buckets may overlap
    num_grad_buffers = 2      # gradients ping-pong between two symmetric buffers (one per graph slot)

    def alloc(self, n, be):
        """Calls in order: weight arena, gradient arena 0, gradient arena 1 (all symmetric)."""
        assert n <= self.n, "arena larger than the communicator's capacity"
        off = (self.off_w, self.off_g, self.off_g2)[self._next]
        self._next += 1
        return self._tensor(off, n)

    def alloc_arenas(self, n, be, ping_pong: bool):
        """([weight buffers], [gradient buffers]) in symmetric memory: two of each when the step is replayed
        from graphs (step t reads W[t & 1], accumulates into G[t & 1]; the update writes W[(t + 1) & 1])."""
        assert n <= self.n, "arena larger than the communicator's capacity"
        bucketed = ping_pong and os.environ.get("B200DIST_BUCKETS", "1") != "0"
        ws = [self._tensor(self.off_w, n)] + ([self._tensor(self.off_w2, n)] if bucketed else [])
        gs = [self._tensor(self.off_g, n)] + ([self._tensor(self.off_g2, n)] if ping_pong else [])
        return ws, gs

    def scatter_target(self, g):
        """Where a weight-gradient GEMM may add its result directly (GEMM -> reduce-scatter, ops/gemm.py):
        gradient arena ``g`` of every replica as mapped in this process."""
        if os.environ.get("B200DIST_FUSED_RS", "1") == "0":
            return None
        off = self._off_of(g)
        return {"bases": [self.base[p] + off for p in range(self.world)], "local": g, "rank": self.rank,
                "world": self.world,
                "min_floats": int(float(os.environ.get("B200DIST_FUSED_RS_MIN_KB", "256")) * 256)}

    def _off_of(self, t):
        return t.data_ptr() - self.base[self.rank]

    def pick(self, n_floats: int) -> int:
        return self.algo if self.algo >= 0 else int(
            self._lib.b2d_allreduce_pick_algo(n_floats, self.world, 1 if self.mc_base else 0))

    def make_step_updates(self, be, w_in, w_out, g, g_other, n_params, lr, early, late_segments):
        """The update of one step as a bucket schedule (engine/trainer.py):

        * ``early`` — list of [begin, end) float ranges, each reduced by its own fused all-reduce+SGD launch on
          the communication branch of the step graph as soon as its gradients are final.  Reads ``w_in`` / ``g``,
          writes ``w_out`` of every replica, no closing barrier (B2D_AR_NO_EXIT_BARRIER).
        * ``late_segments`` — the ranges that are final only when the backward pass ends (first layer + metric
          tail), on the critical path: the LL exchange when they fit, else one fused launch with its barriers.
          Either way this launch closes the step for all buckets.

        Returns (list of early launches, late launch, description)."""
        n_total = w_in.numel()
        fns, names = [], []
        for (b0, e0, scattered) in early:
            fns.append(self.make_update(be, w_in, g, n_params, lr, grads_other=g_other, segments=[(b0, e0)],
                                        channel=0, w_out=w_out, flags=AR_NO_EXIT_BARRIER,
                                        algo=4 if scattered else None))
            names.append("reduce-scattered by the weight-gradient GEMM" if scattered
                         else ALGO_NAMES[self.pick(e0 - b0)])
        n_late = sum(e - b for b, e in late_segments)
        if n_late <= self.ll_max_floats and os.environ.get("B200DIST_LL", "1") != "0":
            d = LlDesc()
            for p in range(self.world):
                d.inbox[p] = self.base[p] + self.off_ll
                d.epoch[p] = self.base[p] + self.off_ll_epoch
                d.w[p] = self.base[p] + self._off_of(w_in)
                d.w_out[p] = self.base[p] + self._off_of(w_out)
                d.g[p] = self.base[p] + self._off_of(g)
                d.g_zero[p] = self.base[p] + self._off_of(g_other)
            d.nseg = len(late_segments)
            for k, (b0, e0) in enumerate(late_segments):
                d.seg_begin[k], d.seg_end[k] = b0, e0
            d.n_params, d.lr, d.rank, d.world, d.emulate = n_params, float(lr), self.rank, self.world, 0
            self._descs = getattr(self, "_descs", []) + [d]
            lib = self._lib

            def late():
                nl.check(lib.b2d_ll_allreduce_sgd(C.byref(d), torch.cuda.current_stream().cuda_stream),
                         "b2d_ll_allreduce_sgd")
            late_name = f"LL push exchange ({n_late} values)"
        else:
            segs = list(late_segments)
            late = self.make_update(be, w_in, g, n_params, lr, grads_other=g_other, segments=segs, channel=1,
                                    w_out=w_out)
            late_name = ALGO_NAMES[self.pick(max(e - b for b, e in segs))]
        desc = (f"{len(early)} overlapped bucket(s) [{', '.join(names)}] + closing bucket [{late_name}]"
                if early else f"closing bucket [{late_name}]")
        return fns, late, desc

    def make_update(self, be, params, grads, n_params, lr, grads_other=None, segments=None, channel=0,
                    algo=None, w_out=None, flags=0):
        """Fused all-reduce+SGD over ``grads``; ``grads_other`` (the other ping-pong buffer of this
        rank) is zeroed by the same kernel, which lets one-shot run with a single barrier.
        ``segments``: up to two [begin, end) float ranges (bucketing); ``channel``: signal pad."""
        d = AllreduceDesc()
        off_g, off_w = self._off_of(grads), self._off_of(params)
        assert off_g in (self.off_g, self.off_g2) and off_w in (self.off_w, self.off_w2)
        for p in range(self.world):
            d.w[p] = self.base[p] + off_w
            d.g[p] = self.base[p] + off_g
            d.sig[p] = self.base[p] + self.off_sig + channel * self.sig_bytes
            d.epoch[p] = self.base[p] + self.off_epoch + channel * self.epoch_bytes
            if w_out is not None:
                d.w_out[p] = self.base[p] + self._off_of(w_out)
        if self.mc_base:
            d.w_mc = self.mc_base + off_w
            d.g_mc = self.mc_base + off_g
            if w_out is not None:
                d.w_out_mc = self.mc_base + self._off_of(w_out)
        d.flags = flags
        d.g_zero = grads_other.data_ptr() if grads_other is not None else None
        n_total = params.numel()
        assert params.data_ptr() == d.w[self.rank] and grads.data_ptr() == d.g[self.rank]
        assert n_total % 4 == 0 and n_params % 4 == 0
        d.n_params, d.n_total, d.lr = n_params, n_total, float(lr)
        d.rank, d.world, d.emulate, d.plain, d.blocks = self.rank, self.world, 0, 0, 0
        d.algo = self.algo if algo is None else algo
        if segments:
            d.nseg = len(segments)
            for k, (b0, e0) in enumerate(segments):
                d.seg_begin[k], d.seg_end[k] = b0, e0
        self._descs = getattr(self, "_descs", []) + [d]
        lib = self._lib

        def run():
            nl.check(lib.b2d_allreduce_sgd(C.byref(d), torch.cuda.current_stream().cuda_stream),
                     "b2d_allreduce_sgd")
        return run

    # -- control plane ------------------------------------------------------------------
    def barrier(self):
        torch.cuda.synchronize(self.device)
        self._dist.barrier()

    def all_equal(self, tensor) -> bool:
        torch.cuda.synchronize(self.device)
        flat = tensor.detach().reshape(-1).float().cpu().contiguous()
        mx, mn = flat.clone(), flat.clone()
        self._dist.all_reduce(mx, op=self._dist.ReduceOp.MAX)
        self._dist.all_reduce(mn, op=self._dist.ReduceOp.MIN)
        return bool(torch.equal(mx, mn))

    def close(self):
        if getattr(self, "_h", None):
            try:
                torch.cuda.synchronize(self.device)
                self._dist.barrier()
            except Exception:
                pass
            self._lib.b2d_symm_destroy(self._h)
            self._h = None
