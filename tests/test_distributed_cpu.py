"""Multi-worker training on the CPU backend, world_size 2 (BASELINE.json config 1):
the README's distributed flow with a localhost TF_CONFIG (README.md:318-393)."""
import json

import numpy as np
import pytest

from distributed_b200.launch import barrier_apply


def _train(ctx, num_workers, epochs=3, steps=5, per_worker=64, export_dir=None, seed_data=0):
    """What each README worker runs: TF_CONFIG -> strategy -> scope -> fit."""
    import os
    os.environ["B200DIST_BACKEND"] = "cpu"
    os.environ["TF_CONFIG"] = json.dumps(ctx.tf_config(base_port=int(ctx.address[0].split(":")[1]) % 20000 + 20000))
    import numpy as np
    import distributed_b200 as tf
    strategy = tf.distribute.experimental.MultiWorkerMirroredStrategy()
    batch_size = per_worker * num_workers
    rng = np.random.default_rng(seed_data)
    x = rng.random((batch_size * steps * epochs, 28, 28, 1))
    y = rng.integers(0, 10, len(x))
    with strategy.scope():
        model = tf.models.mnist_cnn(learning_rate=0.01)
    h = model.fit(x, y, batch_size=batch_size, epochs=epochs, steps_per_epoch=steps, verbose=0)
    w = model.get_weights()
    digest = float(sum(np.abs(a).astype(np.float64).sum() for a in w))
    if export_dir is not None:
        model.save(os.path.join(export_dir, f"trained-{ctx.partition}.hdf5"))
    return json.dumps({"loss": h.history["loss"], "acc": h.history["accuracy"], "digest": digest,
                       "max_acc": repr(max(h.history["accuracy"])), "n": strategy.num_replicas_in_sync})


def test_two_workers_report_identical_metrics_and_weights(tmp_path):
    out = barrier_apply(lambda ctx: _train(ctx, 2, export_dir=str(tmp_path)), 2, timeout=300)
    r = [json.loads(s) for s in out]
    assert r[0]["n"] == 2
    assert r[0]["loss"] == r[1]["loss"] and r[0]["acc"] == r[1]["acc"]       # N ranks, one value (README.md:229-231)
    assert r[0]["max_acc"] == r[1]["max_acc"]
    assert r[0]["digest"] == r[1]["digest"]                                   # mirrored weights, bit for bit
    assert abs(r[0]["loss"][0] - np.log(10)) < 0.15                           # starts near ln 10 (README.md:413)
    for a in r[0]["acc"]:                                                     # accuracy = integer / (5 * 128)
        assert abs(a * 640 - round(a * 640)) < 1e-3
    # every worker saved its own file; they hold the same model (README.md:237-240)
    import distributed_b200 as tf
    w0 = tf.keras.models.load_model(tmp_path / "trained-0.hdf5").get_weights()
    w1 = tf.keras.models.load_model(tmp_path / "trained-1.hdf5").get_weights()
    assert all(np.array_equal(a, b) for a, b in zip(w0, w1))


def test_two_workers_equal_one_worker_with_the_global_batch():
    """N ranks x batch 64 == 1 rank x batch 64N on the same global data order (sync-DP maths)."""
    two = [json.loads(s) for s in barrier_apply(lambda ctx: _train(ctx, 2, epochs=1, steps=3), 2, timeout=300)]
    one = [json.loads(s) for s in barrier_apply(lambda ctx: _train(ctx, 1, epochs=1, steps=3, per_worker=128), 1,
                                                timeout=300)]
    assert abs(two[0]["loss"][0] - one[0]["loss"][0]) < 1e-4
    assert abs(two[0]["digest"] - one[0]["digest"]) < 1e-2 * 1e-3 * one[0]["digest"] + 1e-3


def _shm_allreduce(ctx):
    import numpy as np
    import torch
    from distributed_b200.parallel.comm import ShmComm
    c = ShmComm("t" + ctx.address[0].split(":")[1], ctx.partition, len(ctx.address), 1 << 16)
    out = []
    for n in (1, 7, 1000, 4099):
        t = torch.arange(n, dtype=torch.float32) * (ctx.partition + 1)
        c.allreduce(t)
        out.append(bool(torch.equal(t, torch.arange(n, dtype=torch.float32) * 6)))
    same = c.all_equal(torch.ones(10))
    diff = c.all_equal(torch.full((10,), float(ctx.partition)))
    c.close()
    return json.dumps([out, same, diff])


def test_shm_allreduce_three_ranks_odd_sizes():
    res = [json.loads(s) for s in barrier_apply(_shm_allreduce, 3, timeout=120)]
    for out, same, diff in res:
        assert all(out) and same is True and diff is False


def _shm_after_crash(ctx):
    """Rank 1 starts first and finds the segment a crashed earlier run left under the same name; rank 0 arrives
    late and replaces it.  Rank 1 must drop the stale segment (its creator process is gone) and join the new one."""
    import struct
    import subprocess
    import time
    import torch
    from distributed_b200.parallel.comm import ShmComm
    job = "stale" + ctx.address[0].split(":")[1]
    cap, world = 1 << 12, 2
    path = f"/dev/shm/b200dist-{job}"
    if ctx.partition == 1:
        dead = subprocess.Popen(["true"])
        dead.wait()                                              # a pid that no longer exists
        hdr = struct.pack("<IIIIi4xqi4xq", 0xB200D157, 2, 1, 7, world, cap, dead.pid, 0)
        with open(path, "wb") as f:
            f.write(hdr.ljust(64, b"\0") + b"\0" * (4 * cap * (world + 1)))
    else:
        time.sleep(1.0)                                          # let rank 1 open the stale segment first
    c = ShmComm(job, ctx.partition, world, cap, timeout_s=30.0)
    t = torch.full((100,), float(ctx.partition + 1))
    c.allreduce(t)
    ok = bool((t == 3.0).all())
    c.close()
    return json.dumps(ok)


def test_shm_attach_ignores_the_segment_of_a_crashed_run():
    res = barrier_apply(_shm_after_crash, 2, timeout=120)
    assert [json.loads(s) for s in res] == [True, True], res


def _train_bn(ctx, num_workers, per_worker=8):
    """A BatchNorm model: per-replica batch statistics, moving statistics averaged over replicas."""
    import os
    os.environ["B200DIST_BACKEND"] = "cpu"
    os.environ["TF_CONFIG"] = json.dumps(ctx.tf_config(base_port=int(ctx.address[0].split(":")[1]) % 20000 + 20000))
    import numpy as np
    import distributed_b200 as tf
    L = tf.keras.layers
    strategy = tf.distribute.experimental.MultiWorkerMirroredStrategy()
    rng = np.random.default_rng(5)
    x = rng.standard_normal((per_worker * num_workers, 8, 8, 4)).astype(np.float32)
    x[per_worker:] += 3.0                      # the second replica's shard has a different mean
    y = rng.integers(0, 10, len(x))
    with strategy.scope():
        model = tf.keras.Sequential([L.BatchNormalization(momentum=0.5, input_shape=(8, 8, 4)), L.ReLU(),
                                     L.GlobalAveragePooling2D(), L.Dense(10)])
        model.compile(loss=tf.keras.losses.SparseCategoricalCrossentropy(from_logits=True),
                      optimizer=tf.keras.optimizers.SGD(0.1), metrics=["accuracy"])
    model.fit(x, y, batch_size=len(x), epochs=1, steps_per_epoch=1, verbose=0)
    w = dict(zip(model.weight_names(), model.get_weights()))
    shard_means = [x[r * per_worker:(r + 1) * per_worker].mean((0, 1, 2)).tolist() for r in range(num_workers)]
    return json.dumps({"mm": w["batch_normalization/moving_mean:0"].tolist(),
                       "mv": w["batch_normalization/moving_variance:0"].tolist(),
                       "gamma": w["batch_normalization/gamma:0"].tolist(), "shard_means": shard_means})


def test_batchnorm_moving_statistics_are_averaged_over_replicas():
    r = [json.loads(s) for s in barrier_apply(lambda ctx: _train_bn(ctx, 2), 2, timeout=300)]
    assert r[0]["mm"] == r[1]["mm"] and r[0]["mv"] == r[1]["mv"] and r[0]["gamma"] == r[1]["gamma"]
    # moving_mean = 0.5 * 0 + 0.5 * mean over replicas of the per-replica batch mean
    expect = 0.5 * np.mean(np.asarray(r[0]["shard_means"]), axis=0)
    assert np.allclose(r[0]["mm"], expect, atol=1e-5)


def _train_residual(ctx, num_workers, per_worker):
    """A residual network without BatchNorm (so that N x B and 1 x NB are the same mathematics)."""
    import os
    os.environ["B200DIST_BACKEND"] = "cpu"
    os.environ["TF_CONFIG"] = json.dumps(ctx.tf_config(base_port=int(ctx.address[0].split(":")[1]) % 20000 + 20000))
    import numpy as np
    import distributed_b200 as tf
    L = tf.keras.layers
    strategy = tf.distribute.experimental.MultiWorkerMirroredStrategy()
    rng = np.random.default_rng(9)
    gb = per_worker * num_workers
    x = rng.standard_normal((gb * 3, 10, 10, 3)).astype(np.float32)
    y = rng.integers(0, 4, len(x))
    with strategy.scope():
        model = tf.keras.Sequential([
            L.Conv2D(8, 3, padding="same", activation="relu", input_shape=(10, 10, 3)),
            L.Residual([L.Conv2D(4, 1, activation="relu"), L.Conv2D(8, 3, padding="same")]),
            L.MaxPooling2D(),
            L.Residual([L.Conv2D(4, 1, activation="relu"), L.Conv2D(16, 3, strides=2, padding="same")],
                       [L.Conv2D(16, 1, strides=2, padding="same", use_bias=False)]),
            L.GlobalAveragePooling2D(), L.Dense(4)])
        model.compile(loss=tf.keras.losses.SparseCategoricalCrossentropy(from_logits=True),
                      optimizer=tf.keras.optimizers.SGD(0.05), metrics=["accuracy"])
    h = model.fit(x, y, batch_size=gb, epochs=1, steps_per_epoch=3, verbose=0)
    return json.dumps({"loss": h.history["loss"],
                       "digest": [float(np.abs(a).astype(np.float64).sum()) for a in model.get_weights()]})


def test_residual_network_two_workers_equal_one_worker():
    two = [json.loads(s) for s in barrier_apply(lambda ctx: _train_residual(ctx, 2, 6), 2, timeout=300)]
    one = json.loads(barrier_apply(lambda ctx: _train_residual(ctx, 1, 12), 1, timeout=300)[0])
    assert two[0] == two[1]                                        # mirrored: identical loss and weights
    assert abs(two[0]["loss"][0] - one["loss"][0]) < 1e-4
    for a, b in zip(two[0]["digest"], one["digest"]):
        assert abs(a - b) <= 1e-4 * max(abs(a), 1.0)
