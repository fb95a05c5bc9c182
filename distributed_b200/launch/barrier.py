"""Barrier-style gang launcher — the equivalent of sparklyr's
``spark_apply(f, barrier = TRUE)`` path (README.md:170-224).

``barrier_apply(fn, n)`` starts ``n`` worker processes *together* (one per GPU on the B200
box), hands each a context with

* ``ctx.address``   — list of every task's ``host:port`` (``barrier$address``, README.md:181)
* ``ctx.partition`` — this task's 0-based index       (``barrier$partition``, README.md:182, 240)

and collects exactly one string per task, in partition order.  An exception inside a task
does not fail the job: its message comes back as that task's string, which is what the
README's ``tryCatch(..., error = function(e) e$message)`` does (README.md:176, 221).  If a
task dies without reporting (killed, segfault) or the gang exceeds ``timeout``, every
remaining task is terminated — fail-stop, like a Spark barrier stage.
"""
from __future__ import annotations

import multiprocessing as mp
import os
import socket
import time
import traceback
from dataclasses import dataclass
from typing import Callable, Sequence


@dataclass
class BarrierContext:
    address: list[str]
    partition: int

    @property
    def num_tasks(self) -> int:
        return len(self.address)

    def tf_config(self, base_port: int = 8000) -> dict:
        """TF_CONFIG built the way the README does it (README.md:180-183)."""
        from ..parallel.cluster import ClusterSpec
        return ClusterSpec.from_barrier(self.address, self.partition, base_port).to_dict()


class BarrierResult(list):
    """One string per task; ``result['address']`` mimics the one-column tibble (README.md:226-231)."""

    def __init__(self, values, column="address"):
        super().__init__(values)
        self.column = column

    def __getitem__(self, k):
        if isinstance(k, str):
            if k != self.column:
                raise KeyError(k)
            return list(self)
        return super().__getitem__(k)

    def to_table(self) -> str:
        w = max([len(self.column)] + [len(repr(v)) for v in self]) + 2
        lines = [f"# A table: {len(self)} x 1", f"  {self.column}", "  <chr>"]
        lines += [f"{i + 1} {v}" for i, v in enumerate(self)]
        return "\n".join(lines)


class BarrierError(RuntimeError):
    pass


def _free_ports(n: int, host: str) -> list[int]:
    socks, ports = [], []
    for _ in range(n):
        s = socket.socket()
        s.bind((host, 0))
        socks.append(s)
        ports.append(s.getsockname()[1])
    for s in socks:
        s.close()
    return ports


def _worker(payload: bytes, ctx: BarrierContext, conn, env: dict, capture_errors: bool):
    try:
        os.environ.update(env)
        import cloudpickle
        fn, args = cloudpickle.loads(payload)
        try:
            out = fn(*args, ctx) if args else fn(ctx)
            conn.send(("ok", "" if out is None else str(out)))
        except BaseException as e:  # noqa: BLE001 - the task's error text IS the result
            if not capture_errors:
                raise
            msg = str(e) or type(e).__name__
            conn.send(("err", msg))
    except BaseException:  # noqa: BLE001
        try:
            conn.send(("crash", traceback.format_exc()))
        except Exception:
            pass
    finally:
        conn.close()


def barrier_apply(fn: Callable, n: int, *, args: Sequence | None = None, host: str = "127.0.0.1",
                  timeout: float = 600.0, column: str = "address", capture_errors: bool = True,
                  env: dict | None = None, assign_gpus: bool = True) -> BarrierResult:
    """Run ``fn(ctx)`` (or ``fn(partition_arg, ctx)`` when ``args`` is given) on ``n`` gang-
    scheduled worker processes and return their ``n`` result strings in partition order."""
    import cloudpickle
    if n < 1:
        raise ValueError("n must be >= 1")
    if args is not None and len(args) != n:
        raise ValueError("args must have one entry per task")
    ports = _free_ports(n, host)
    address = [f"{host}:{p}" for p in ports]
    mpctx = mp.get_context("spawn")
    procs, conns = [], []
    for i in range(n):
        parent, child = mpctx.Pipe(duplex=False)
        wenv = dict(env or {})
        if assign_gpus:
            wenv.setdefault("LOCAL_RANK", str(i))
        payload = cloudpickle.dumps((fn, (args[i],) if args is not None else ()))
        p = mpctx.Process(target=_worker, args=(payload, BarrierContext(list(address), i), child, wenv,
                                                capture_errors), daemon=True)
        procs.append(p)
        conns.append(parent)
    for p in procs:          # the gang starts together: all or nothing
        p.start()
    results: list[str | None] = [None] * n
    crashed: list[str] = []
    deadline = time.monotonic() + timeout
    pending = set(range(n))
    try:
        while pending:
            for i in list(pending):
                if conns[i].poll(0.05):
                    try:
                        kind, text = conns[i].recv()
                    except EOFError:
                        kind, text = "crash", f"task {i} closed its pipe without a result"
                    pending.discard(i)
                    if kind == "crash":
                        crashed.append(f"task {i}: {text}")
                    else:
                        results[i] = text
                elif not procs[i].is_alive() and not conns[i].poll(0):
                    pending.discard(i)
                    crashed.append(f"task {i} exited with code {procs[i].exitcode} without a result")
            if crashed:
                raise BarrierError("barrier stage failed; killing the gang:\n" + "\n".join(crashed))
            if time.monotonic() > deadline:
                raise BarrierError(f"barrier stage timed out after {timeout:.0f} s; "
                                   f"tasks {sorted(pending)} did not finish")
    finally:
        for p in procs:
            if p.is_alive() and (pending or crashed):
                p.terminate()
        for p in procs:
            p.join(timeout=10)
            if p.is_alive():
                p.kill()
    return BarrierResult([r if r is not None else "" for r in results], column)
