"""Multi-GPU tier (needs >= 2 B200s): the README's distributed flow with one process per GPU,
gradients reduced by the fused peer-memory kernels over NVLink (no NCCL on the path)."""
import json

import numpy as np
import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _worker(ctx, n, algo, steps=5, epochs=2, env=None):
    import os
    os.environ.update(env or {})
    os.environ["B200DIST_ALLREDUCE"] = algo
    os.environ["TF_CONFIG"] = json.dumps(ctx.tf_config(base_port=int(ctx.address[0].split(":")[1]) % 20000 + 30000))
    import numpy as np
    import torch
    import distributed_b200 as tf
    strategy = tf.distribute.experimental.MultiWorkerMirroredStrategy()
    rng = np.random.default_rng(0)
    gb = 64 * n
    x = rng.random((gb * min(steps * epochs, 8), 28, 28, 1)).astype(np.float32)     # longer runs wrap around
    y = rng.integers(0, 10, len(x))
    with strategy.scope():
        model = tf.models.mnist_cnn(learning_rate=0.01)
    try:
        h = model.fit(x, y, batch_size=gb, epochs=epochs, steps_per_epoch=steps, verbose=0)
    except RuntimeError as e:
        if "multicast is not available" in str(e):
            return json.dumps({"skip": str(e)})
        raise
    w = model.get_weights()
    t = model._trainer
    out = {"loss": h.history["loss"], "acc": h.history["accuracy"],
           "digest": [float(np.abs(a).astype(np.float64).sum()) for a in w],
           "device": str(t.be.device), "comm": t.comm.description, "graph": t.use_graph,
           "kernels": t.kernels_per_step, "schedule": getattr(t, "update_schedule", "single fused launch")}
    t.close()
    return json.dumps(out)


_BASE = {}


def _baseline(n, algo, steps):
    """Default-schedule run, shared by the tests that compare against it (one gang launch instead of one each)."""
    from distributed_b200.launch import barrier_apply
    key = (n, algo, steps)
    if key not in _BASE:
        _BASE[key] = json.loads(barrier_apply(lambda ctx: _worker(ctx, n, algo, steps=steps, epochs=1), n,
                                              timeout=900)[0])
    return _BASE[key]


def _ngpu():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("algo", ["oneshot", "twoshot", "nvls", "pull", "auto"])
def test_replicas_stay_identical(algo):
    from distributed_b200.launch import barrier_apply
    n = min(_ngpu(), 8)
    res = [json.loads(s) if s.startswith("{") else s for s in
           barrier_apply(lambda ctx: _worker(ctx, n, algo), n, timeout=600)]
    assert all(isinstance(r, dict) for r in res), res
    if "skip" in res[0]:
        pytest.skip(res[0]["skip"])
    assert len({r["device"] for r in res}) == n and all(r["graph"] for r in res)
    for r in res[1:]:
        assert r["loss"] == res[0]["loss"] and r["acc"] == res[0]["acc"]        # N ranks, one value
        assert r["digest"] == res[0]["digest"]                                    # mirrored, bit for bit
    assert abs(res[0]["loss"][0] - np.log(10)) < 0.15
    assert "NVLink" in res[0]["comm"]


def test_n_gpus_match_one_gpu_with_the_global_batch():
    from distributed_b200.launch import barrier_apply
    import distributed_b200 as tf
    n = min(_ngpu(), 8)
    multi = json.loads(barrier_apply(lambda ctx: _worker(ctx, n, "auto", steps=3, epochs=1), n, timeout=600)[0])
    rng = np.random.default_rng(0)
    gb = 64 * n
    x = rng.random((gb * 3, 28, 28, 1)).astype(np.float32)
    y = rng.integers(0, 10, len(x))
    m = tf.models.mnist_cnn(learning_rate=0.01)
    h = m.fit(x, y, batch_size=gb, epochs=1, steps_per_epoch=3, verbose=0)
    assert abs(h.history["loss"][0] - multi["loss"][0]) < 2e-3
    single = [float(np.abs(a).astype(np.float64).sum()) for a in m.get_weights()]
    for a, b in zip(single, multi["digest"]):
        assert abs(a - b) <= 2e-3 * max(abs(a), 1.0)


@pytest.mark.parametrize("env", [{"B200DIST_BUCKETS": "0"}, {"B200DIST_LL": "0"}, {"B200DIST_SIDE_WGRAD": "0"},
                                 {"B200DIST_AR_STRICT": "1"}, {"B200DIST_BUCKET_MB": "0.001"}])
def test_optional_update_schedules_keep_replicas_identical_and_match_the_default(env):
    """The default update is a bucket schedule overlapped with the backward pass (weight ping-pong, early buckets
    on the communication branch, LL closing exchange).  One fused launch (BUCKETS=0), a barrier-based closing
    bucket (LL=0), a single-chain step, release/acquire entry barriers and one bucket per layer are different
    graphs over the same arithmetic: replicas must stay bit-identical and the trajectory must match the default."""
    from distributed_b200.launch import barrier_apply
    n = min(_ngpu(), 8)
    base = _baseline(n, "auto", 4)
    assert "overlapped bucket" in base["schedule"]
    res = [json.loads(s) if s.startswith("{") else s for s in
           barrier_apply(lambda ctx: _worker(ctx, n, "auto", steps=4, epochs=1, env=env), n, timeout=600)]
    assert all(isinstance(r, dict) for r in res), res
    for r in res[1:]:
        assert r["loss"] == res[0]["loss"] and r["digest"] == res[0]["digest"]
    assert abs(res[0]["loss"][0] - base["loss"][0]) < 1e-3
    for a, b in zip(res[0]["digest"], base["digest"]):
        assert abs(a - b) <= 1e-3 * max(abs(a), 1.0)


@pytest.mark.parametrize("algo", ["auto", "pull", "twoshot"])
def test_debug_mode_stress_random_delays_poison_and_monotone_flags(algo):
    """the race detector of the peer-memory path.  B200DIST_AR_DEBUG=1 pauses pseudo-random (rank,
    block, epoch) combinations at every phase of the all-reduce kernels, poisons each consumed gradient buffer
    with NaNs for 20 us before zeroing it, and traps when a barrier flag is not monotone.  A read outside the
    protocol's windows turns the weights into NaNs or makes replicas diverge.  The arithmetic is unchanged, so the
    trajectory must match the run without the debug mode (up to the summation order of the atomically accumulated
    weight gradients, which differs from run to run)."""
    from distributed_b200.launch import barrier_apply
    n = min(_ngpu(), 8)
    steps = 300
    ref = _baseline(n, algo, steps)
    res = [json.loads(s) if s.startswith("{") else s for s in
           barrier_apply(lambda ctx: _worker(ctx, n, algo, steps=steps, epochs=1,
                                             env={"B200DIST_AR_DEBUG": "1"}), n, timeout=900)]
    assert all(isinstance(r, dict) for r in res), res
    for r in res:
        assert all(np.isfinite(v) for v in r["digest"]) and np.isfinite(r["loss"][0])
        assert r["digest"] == res[0]["digest"]
    assert abs(res[0]["loss"][0] - ref["loss"][0]) < 2e-3
    for a, b in zip(res[0]["digest"], ref["digest"]):
        assert abs(a - b) <= 5e-3 * max(abs(a), 1.0)        # sum |w| per tensor; a wrong update is off by far more


def _allreduce_vs_nccl(ctx, n):
    """Plain all-reduce of the peer-memory kernels against torch.distributed.all_reduce (NCCL), SURVEY 4.2."""
    import os
    import ctypes as C
    os.environ["TF_CONFIG"] = json.dumps(ctx.tf_config(base_port=int(ctx.address[0].split(":")[1]) % 20000 + 30000))
    import torch
    import torch.distributed as dist
    import distributed_b200 as tf
    from distributed_b200._native import cuda_lib as nl
    from distributed_b200.parallel import symm
    strategy = tf.distribute.experimental.MultiWorkerMirroredStrategy()
    strategy._ensure_group()
    rank = strategy.rank
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    nccl = dist.new_group(backend="nccl")
    sizes = [4, 4 * (n + 1), 4 * (3 * n + 1), 1000 * 4 + 4, 347172, (1 << 20) + 4 * (n - 1), (1 << 22) + 12]
    comm = symm.SymmComm(strategy, f"cmp-{ctx.address[0]}", max(sizes) + 64, dev)
    lib = nl.lib()
    out = {"sizes": sizes, "max_err": [], "digests": []}
    algos = [0, 1] + ([2] if comm.mc_base else [])
    for n_floats in sizes:
        g = torch.Generator(device=dev).manual_seed(1000 * rank + n_floats)
        src = torch.randn(n_floats, device=dev, generator=g)
        ref = src.clone()
        dist.all_reduce(ref, group=nccl)
        for algo in algos:
            G = comm._tensor(comm.off_g, n_floats)
            O = comm._tensor(comm.off_w, n_floats)
            G.copy_(src)
            O.zero_()
            comm.barrier()
            d = symm.AllreduceDesc()
            for p in range(n):
                d.w[p] = d.out[p] = comm.base[p] + comm.off_w
                d.g[p] = comm.base[p] + comm.off_g
                d.sig[p], d.epoch[p] = comm.base[p] + comm.off_sig, comm.base[p] + comm.off_epoch
            if comm.mc_base:
                d.w_mc, d.g_mc = comm.mc_base + comm.off_w, comm.mc_base + comm.off_g
            d.n_params = d.n_total = n_floats
            d.lr, d.rank, d.world, d.algo, d.emulate, d.plain, d.blocks = 0.0, rank, n, algo, 0, 1, 0
            nl.check(lib.b2d_allreduce_sgd(C.byref(d), torch.cuda.current_stream().cuda_stream), "allreduce")
            comm.barrier()
            res = O if algo == 0 else G          # one-shot writes `out`; two-shot / NVLS reduce in place
            out["max_err"].append(float((res - ref).abs().max() / ref.abs().max()))
            out["digests"].append(res.double().sum().item().hex() + res.view(torch.int32).sum().item().__repr__())
    comm.close()
    return json.dumps(out)


def test_plain_allreduce_matches_nccl_on_odd_sizes_and_is_bitwise_identical_across_ranks():
    from distributed_b200.launch import barrier_apply
    n = min(_ngpu(), 8)
    res = [json.loads(s) if s.startswith("{") else s for s in
           barrier_apply(lambda ctx: _allreduce_vs_nccl(ctx, n), n, timeout=600)]
    assert all(isinstance(r, dict) for r in res), res
    assert max(res[0]["max_err"]) < 1e-5, res[0]["max_err"]                 # vs NCCL (its summation order differs)
    for r in res[1:]:
        assert r["digests"] == res[0]["digests"]                              # every rank holds the same bits
