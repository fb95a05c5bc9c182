"""barrier_apply: the spark_apply(barrier = TRUE) contract (README.md:175-183, 221, 226-231)."""
import pytest

from distributed_b200.launch import BarrierError, barrier_apply, sdf_len, spark_connect


def _echo(ctx):
    return f"{ctx.partition}/{len(ctx.address)}/{ctx.address[ctx.partition]}"


def test_returns_one_string_per_task_in_partition_order():
    out = barrier_apply(_echo, 3, timeout=120)
    assert len(out) == 3 and out["address"] == list(out)
    for i, s in enumerate(out):
        part, n, addr = s.split("/")
        assert int(part) == i and int(n) == 3 and addr.startswith("127.0.0.1:")
    assert len({s.split("/")[2] for s in out}) == 3          # every task has its own port
    assert "3 x 1" in out.to_table()


def _fail_on_one(ctx):
    if ctx.partition == 1:
        raise ValueError("boom in task 1")
    return "fine"


def test_exception_comes_back_as_message_string():
    out = barrier_apply(_fail_on_one, 3, timeout=120)          # tryCatch(..., error = e$message)
    assert list(out) == ["fine", "boom in task 1", "fine"]


def _die(ctx):
    import os
    if ctx.partition == 0:
        os._exit(7)
    import time
    time.sleep(60)


def test_dead_task_kills_the_gang():
    with pytest.raises(BarrierError):
        barrier_apply(_die, 2, timeout=120)


def test_tf_config_from_barrier_context():
    def fn(ctx):
        import json
        return json.dumps(ctx.tf_config())
    import json
    out = barrier_apply(fn, 2, timeout=120)
    cfgs = [json.loads(s) for s in out]
    assert [c["task"]["index"] for c in cfgs] == [0, 1]
    assert cfgs[0]["cluster"]["worker"] == ["127.0.0.1:8001", "127.0.0.1:8002"]


def test_sparklyr_shaped_flow():
    sc = spark_connect(master="local", config={"spark.dynamicAllocation.enabled": False,
                                               "sparklyr.shell.executor-cores": 8,
                                               "sparklyr.shell.num-executors": 3})
    res = sdf_len(sc, 3, repartition=3).spark_apply(lambda df, barrier: f"{barrier.partition}:{df}",
                                                    barrier=True, columns={"address": "character"},
                                                    timeout=120).collect()
    assert list(res) == ["0:[0]", "1:[1]", "2:[2]"]
    with pytest.raises(ValueError):
        spark_connect(master="yarn")
