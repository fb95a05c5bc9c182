"""MultiWorkerMirroredStrategy: synchronous data parallelism over the workers of TF_CONFIG.

Mirrors ``tf.distribute.experimental.MultiWorkerMirroredStrategy()`` as the README uses it
(README.md:364, 122, 187): a no-argument constructor that reads ``TF_CONFIG``, a ``scope()``
inside which the model and its ``compile`` are created (README.md:375-386), and an implicit
all-reduce of the gradients in ``fit``.  Worker *i* of the cluster spec is replica *i*; on the
B200 box replica *i* drives ``cuda:i`` (one process per GPU).

What is different underneath: variables are slices of one flat arena that is created
identically on every replica (same seed, verified bit-for-bit), and the per-step
"batch_all_reduce: 6 all-reduces" (README.md:403) is one fused kernel over NVLink peer
memory that also averages, applies SGD and aggregates the metrics.
"""
from __future__ import annotations

import contextlib
import enum
import logging
import os
import threading

import torch

from .cluster import ClusterSpec

log = logging.getLogger("distributed_b200")
if not log.handlers:
    _h = logging.StreamHandler()
    _h.setFormatter(logging.Formatter("%(levelname)s:distributed_b200:%(message)s"))
    log.addHandler(_h)
    log.setLevel(os.environ.get("B200DIST_LOG", "INFO"))
    log.propagate = False


class CollectiveCommunication(enum.Enum):
    """Which gradient all-reduce implementation to use (README.md:398 shows AUTO)."""
    AUTO = "AUTO"        # CUDA: fused peer-memory kernels; CPU: shared memory or gloo
    RING = "RING"        # accepted for API compatibility; NVSwitch needs no ring -> same as AUTO
    NCCL = "NCCL"        # torch.distributed NCCL all_reduce + separate SGD kernel (the baseline)
    ONE_SHOT = "ONE_SHOT"
    TWO_SHOT = "TWO_SHOT"
    NVLS = "NVLS"


_state = threading.local()
_global_seed = [1234]


def set_seed(seed: int) -> None:
    """Seed used to initialise mirrored variables (must be equal on every worker)."""
    _global_seed[0] = int(seed)


def get_strategy():
    s = getattr(_state, "stack", None)
    return s[-1] if s else _default_strategy()


_default = None


def _default_strategy():
    global _default
    if _default is None:
        _default = OneDeviceStrategy()
    return _default


class _StrategyBase:
    num_workers = 1
    rank = 0

    @property
    def num_replicas_in_sync(self) -> int:
        return self.num_workers

    @contextlib.contextmanager
    def scope(self):
        stack = getattr(_state, "stack", None)
        if stack is None:
            stack = _state.stack = []
        stack.append(self)
        try:
            yield self
        finally:
            stack.pop()

    def _pick_backend(self, backend):
        name = backend or os.environ.get("B200DIST_BACKEND")
        if name is None:
            name = "cuda" if torch.cuda.is_available() else "cpu"
        if name not in ("cuda", "cpu"):
            raise ValueError("backend must be 'cuda' or 'cpu'")
        if name == "cuda" and not torch.cuda.is_available():
            raise RuntimeError("backend 'cuda' requested but no GPU is visible")
        return name

    def _pick_device(self, rank):
        if self.backend_name != "cuda":
            return torch.device("cpu")
        n = torch.cuda.device_count()
        local = int(os.environ.get("LOCAL_RANK", rank % n))
        return torch.device("cuda", local % n)


class OneDeviceStrategy(_StrategyBase):
    """The implicit strategy outside any scope: one replica, no communication (README 'Local')."""

    def __init__(self, backend: str | None = None):
        self.backend_name = self._pick_backend(backend)
        self.device = self._pick_device(0)
        self.cluster = ClusterSpec()
        self.seed = _global_seed[0]
        self.communication = CollectiveCommunication.AUTO

    def make_comm(self, capacity, be):
        from .comm import SingleComm
        return SingleComm()


class MultiWorkerMirroredStrategy(_StrategyBase):
    _instances = 0

    def __init__(self, communication: CollectiveCommunication = CollectiveCommunication.AUTO,
                 cluster_resolver=None, backend: str | None = None):
        self.cluster = cluster_resolver if isinstance(cluster_resolver, ClusterSpec) else ClusterSpec.from_env()
        self.num_workers = self.cluster.num_workers
        self.rank = self.cluster.task_index
        self.backend_name = self._pick_backend(backend)
        self.device = self._pick_device(self.rank)
        self.communication = communication
        self.seed = _global_seed[0]
        self._comm_seq = 0
        self._group_ready = False
        MultiWorkerMirroredStrategy._instances += 1
        self._announce()

    # -- logging that mirrors what the TF log tells the user (README.md:395-399) --------
    def _announce(self):
        spec = {"worker": list(self.cluster.workers)}
        log.info("Running independent worker: cluster_spec = %s, task_type = 'worker', task_id = %d, "
                 "control_plane = 'tcp-store'", spec, self.rank)
        dev = str(self.device) if self.backend_name == "cuda" else "/job:worker/task:%d (cpu)" % self.rank
        log.info("Multi-worker mirrored strategy with cluster_spec = %s, task_id = %d, num_workers = %d, "
                 "local_devices = ('%s',), communication = CollectiveCommunication.%s",
                 spec, self.rank, self.num_workers, dev, self.communication.name)

    # -- control plane -----------------------------------------------------------------
    def _ensure_group(self):
        """torch.distributed (gloo) process group used as the control plane and CPU fallback."""
        import torch.distributed as dist
        if self._group_ready or self.num_workers == 1:
            return
        if not dist.is_initialized():
            import datetime
            timeout = datetime.timedelta(seconds=float(os.environ.get("B200DIST_RENDEZVOUS_TIMEOUT", "300")))
            if self.cluster.source == "torchrun-env":
                dist.init_process_group("gloo", timeout=timeout)      # env:// via the launcher's store
            else:
                host0 = self.cluster.host(0)
                dist.init_process_group("gloo", init_method=f"tcp://{host0}:{self.cluster.port(0)}",
                                        rank=self.rank, world_size=self.num_workers, timeout=timeout)
        self._group_ready = True

    def make_comm(self, capacity: int, be):
        from . import comm as C
        if self.num_workers == 1:
            return C.SingleComm()
        self._comm_seq += 1
        tag = f"{self.cluster.job_id()}-{self._comm_seq}"
        if self.backend_name == "cpu":
            if self.cluster.single_host() and self.communication != CollectiveCommunication.RING:
                c = C.ShmComm(tag, self.rank, self.num_workers, max(capacity, 1 << 16))
            else:
                self._ensure_group()
                c = C.TorchDistComm()
        else:
            self._ensure_group()
            if self.communication == CollectiveCommunication.NCCL:
                import torch.distributed as dist
                group = dist.new_group(backend="nccl")
                c = C.TorchDistComm(group)
            else:
                from .symm import SymmComm
                c = SymmComm(self, tag, capacity, be.device)
        log.info("Collective batch_all_reduce: gradients + metrics fused with averaging and SGD, "
                 "num_workers = %d, implementation = %s", self.num_workers, c.description)
        return c


class _Experimental:
    MultiWorkerMirroredStrategy = MultiWorkerMirroredStrategy
    CollectiveCommunication = CollectiveCommunication
