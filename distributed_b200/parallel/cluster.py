"""Cluster specification: the ``TF_CONFIG`` contract of the reference.

The README configures every worker with a JSON environment variable that differs only in
``task.index`` (README.md:322-327; R form README.md:84-89)::

    {"cluster": {"worker": ["172.17.0.6:10090", ...]}, "task": {"type": "worker", "index": 0}}

``ClusterSpec`` parses exactly that, and also synthesises it from a barrier context the way
the sparklyr example does (strip the port of every address, assign port ``8000 + i`` for
i = 1..n, index = partition; README.md:180-183) or from ``torchrun``-style environment
variables (RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT), so one strategy object works under
either launcher.
"""
from __future__ import annotations

import json
import os
import re
import socket
from dataclasses import dataclass, field
from typing import Mapping, Sequence


class ClusterSpecError(ValueError):
    pass


@dataclass(frozen=True)
class ClusterSpec:
    workers: tuple[str, ...] = ("localhost:0",)
    task_type: str = "worker"
    task_index: int = 0
    source: str = field(default="default", compare=False)

    # ---- construction -------------------------------------------------------------
    @staticmethod
    def from_dict(cfg: Mapping, source: str = "dict") -> "ClusterSpec":
        try:
            cluster = cfg["cluster"]
            task = cfg["task"]
        except (KeyError, TypeError) as e:
            raise ClusterSpecError(f"TF_CONFIG must contain 'cluster' and 'task': {e}") from None
        if not isinstance(cluster, Mapping) or "worker" not in cluster:
            raise ClusterSpecError("TF_CONFIG['cluster'] must contain a 'worker' list")
        workers = cluster["worker"]
        if isinstance(workers, str):          # jsonlite without auto_unbox on a 1-worker list
            workers = [workers]
        if not isinstance(workers, Sequence) or len(workers) == 0:
            raise ClusterSpecError("TF_CONFIG['cluster']['worker'] must be a non-empty list")
        for w in workers:
            if not isinstance(w, str) or not re.fullmatch(r"[^:\s]+:\d+", w):
                raise ClusterSpecError(f"worker address {w!r} is not 'host:port'")
        ttype = task.get("type", "worker") if isinstance(task, Mapping) else None
        index = task.get("index") if isinstance(task, Mapping) else None
        if isinstance(ttype, list) and len(ttype) == 1:   # R toJSON without auto_unbox
            ttype = ttype[0]
        if isinstance(index, list) and len(index) == 1:
            index = index[0]
        if ttype != "worker":
            raise ClusterSpecError(f"only task type 'worker' is supported (got {ttype!r}); the "
                                   "reference runs every process as an independent worker")
        if not isinstance(index, int) or isinstance(index, bool):
            raise ClusterSpecError(f"task index must be an integer (got {index!r})")
        if not 0 <= index < len(workers):
            raise ClusterSpecError(f"task index {index} out of range for {len(workers)} workers")
        return ClusterSpec(tuple(workers), "worker", index, source)

    @staticmethod
    def from_json(text: str, source: str = "json") -> "ClusterSpec":
        try:
            cfg = json.loads(text)
        except json.JSONDecodeError as e:
            raise ClusterSpecError(f"TF_CONFIG is not valid JSON: {e}") from None
        return ClusterSpec.from_dict(cfg, source)

    @staticmethod
    def from_env(env: Mapping[str, str] | None = None) -> "ClusterSpec":
        """TF_CONFIG if set; else torchrun-style variables; else a single local worker."""
        env = os.environ if env is None else env
        if env.get("TF_CONFIG"):
            return ClusterSpec.from_json(env["TF_CONFIG"], "TF_CONFIG")
        if "RANK" in env and "WORLD_SIZE" in env:
            world, rank = int(env["WORLD_SIZE"]), int(env["RANK"])
            host = env.get("MASTER_ADDR", "127.0.0.1")
            port = int(env.get("MASTER_PORT", "29500"))
            # torchrun publishes one rendezvous endpoint; derive one port per worker from it
            # (worker 0's address doubles as the control-plane store, as in TF_CONFIG).
            workers = tuple(f"{host}:{port + i}" for i in range(world))
            return ClusterSpec(workers, "worker", rank, "torchrun-env")
        return ClusterSpec()

    @staticmethod
    def from_barrier(addresses: Sequence[str], partition: int, base_port: int = 8000) -> "ClusterSpec":
        """The sparklyr recipe: hosts from ``barrier$address``, ports ``base_port + 1..n``."""
        hosts = [re.sub(r":[0-9]+$", "", a) for a in addresses]
        workers = tuple(f"{h}:{base_port + i}" for i, h in enumerate(hosts, start=1))
        return ClusterSpec.from_dict(
            {"cluster": {"worker": list(workers)}, "task": {"type": "worker", "index": int(partition)}},
            "barrier")

    # ---- views --------------------------------------------------------------------
    @property
    def num_workers(self) -> int:
        return len(self.workers)

    @property
    def is_chief(self) -> bool:
        return self.task_index == 0

    def host(self, i: int | None = None) -> str:
        return self.workers[self.task_index if i is None else i].rsplit(":", 1)[0]

    def port(self, i: int | None = None) -> int:
        return int(self.workers[self.task_index if i is None else i].rsplit(":", 1)[1])

    def to_dict(self) -> dict:
        return {"cluster": {"worker": list(self.workers)},
                "task": {"type": self.task_type, "index": self.task_index}}

    def to_json(self) -> str:
        return json.dumps(self.to_dict())

    def export(self, env: dict | None = None) -> None:
        (os.environ if env is None else env)["TF_CONFIG"] = self.to_json()

    def single_host(self) -> bool:
        """True when every worker address names this machine (one 8xB200 box, or localhost)."""
        hosts = {self.host(i) for i in range(self.num_workers)}
        if len(hosts) == 1:
            return True
        local = {"localhost", "127.0.0.1", socket.gethostname()}
        return hosts <= local

    def job_id(self) -> str:
        """Stable identifier shared by all workers of this cluster (names shm / sockets)."""
        import hashlib
        return hashlib.sha1(",".join(self.workers).encode()).hexdigest()[:12]


def local_cluster(n: int, base_port: int = 0, host: str = "127.0.0.1") -> list[ClusterSpec]:
    """n worker specs on this machine with free ports (multi-process runs without a cluster)."""
    ports = []
    socks = []
    for i in range(n):
        if base_port:
            ports.append(base_port + i)
            continue
        s = socket.socket()
        s.bind((host, 0))
        socks.append(s)
        ports.append(s.getsockname()[1])
    for s in socks:
        s.close()
    workers = tuple(f"{host}:{p}" for p in ports)
    return [ClusterSpec(workers, "worker", i, "local_cluster") for i in range(n)]


def host_ip() -> str:
    """The README's IP discovery helper (README.md:271-275)."""
    try:
        return socket.gethostbyname(socket.gethostname())
    except OSError:
        return "127.0.0.1"
