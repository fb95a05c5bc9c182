"""A tiny stand-in for the sparklyr calls the README makes around the barrier stage
(README.md:172-175, 222-223), so the R recipe translates line by line::

    sc = spark_connect(master="local", config={"sparklyr.shell.num-executors": 3,
                                               "spark.dynamicAllocation.enabled": False})
    result = sdf_len(sc, 3, repartition=3).spark_apply(train, barrier=True,
                                                       columns={"address": "character"}).collect()

There is no Spark here: ``master`` must be ``local`` (one B200 box), the executors are worker
processes, and dynamic allocation must be off because a barrier gang needs a fixed size
(README.md:172 sets ``spark.dynamicAllocation.enabled = FALSE`` for the same reason).
"""
from __future__ import annotations

from dataclasses import dataclass, field

from .barrier import BarrierResult, barrier_apply


@dataclass
class SparkLikeConnection:
    master: str = "local"
    config: dict = field(default_factory=dict)

    @property
    def num_executors(self) -> int:
        return int(self.config.get("sparklyr.shell.num-executors", self.config.get("num_executors", 1)))


def spark_connect(master: str = "local", spark_home=None, config: dict | None = None) -> SparkLikeConnection:
    config = dict(config or {})
    if not master.startswith("local"):
        raise ValueError("only master='local' is available: workers are processes on this box "
                         "(the reference's YARN cluster, README.md:172, is out of scope)")
    if config.get("spark.dynamicAllocation.enabled", False):
        raise ValueError("barrier execution needs spark.dynamicAllocation.enabled = False")
    return SparkLikeConnection(master, config)


@dataclass
class _Frame:
    sc: SparkLikeConnection
    rows: int
    partitions: int

    def spark_apply(self, fn, barrier: bool = False, columns=None, **kw):
        if not barrier:
            raise ValueError("only barrier=True is supported (gang-scheduled tasks)")
        if self.partitions > max(self.sc.num_executors, 1) and "sparklyr.shell.num-executors" in self.sc.config:
            raise ValueError(f"{self.partitions} barrier tasks need {self.partitions} executors, "
                             f"connection has {self.sc.num_executors}")
        column = next(iter(columns)) if columns else "address"
        per = [list(range(self.rows))[i::self.partitions] for i in range(self.partitions)]
        return _Lazy(lambda: barrier_apply(fn, self.partitions, args=per, column=column, **kw))


@dataclass
class _Lazy:
    thunk: object

    def collect(self) -> BarrierResult:
        return self.thunk()


def sdf_len(sc: SparkLikeConnection, n: int, repartition: int | None = None) -> _Frame:
    return _Frame(sc, int(n), int(repartition or 1))
