"""The bucket schedule of the overlapped update (engine/trainer.py::_bucket_plan) on the CPU lowering: which
gradient ranges are reduced when, what rides in the closing exchange, and how GEMM-reduce-scattered ranges split a
bucket (BASELINE.json north star: "bucketed and overlapped with the backward of earlier layers")."""
import torch

import distributed_b200 as tf
from distributed_b200.engine.arena import ArenaLayout
from distributed_b200.engine.cpu_backend import CpuBackend
from distributed_b200.engine.lowering import lower
from distributed_b200.engine.trainer import Trainer


class _Comm:
    ll_max_floats = 16384


def _plan(model, shape, scattered=(), bucket_mb=None, monkeypatch=None):
    lay = ArenaLayout(model.layers)
    P, G = torch.zeros(lay.n_total), torch.zeros(lay.n_total)
    x, y = torch.zeros((2, *shape)), torch.zeros(2, dtype=torch.int32)
    prog = lower(model.layers, 2, CpuBackend(), lay, P, G, x, y, lr=0.01)
    prog.scattered = sorted(scattered)
    fake = type("T", (), {})()
    fake.layout, fake.comm = lay, _Comm()
    if bucket_mb is not None:
        monkeypatch.setenv("B200DIST_BUCKET_MB", str(bucket_mb))
    return lay, prog, Trainer._bucket_plan(fake, prog)


def _covers_everything_once(lay, early, late):
    pieces = sorted([(b, e) for _, b, e, _ in early] + list(late))
    assert pieces[0][0] == 0 and pieces[-1][1] == lay.n_total
    for (b0, e0), (b1, e1) in zip(pieces, pieces[1:]):
        assert e0 == b1, (pieces,)


def test_readme_model_one_overlapped_bucket_and_a_small_closing_exchange():
    lay, prog, (early, late) = _plan(tf.models.mnist_cnn(), (28, 28, 1))
    # grad_final: conv is final only when the backward pass ends; dense-64 and dense-10 earlier
    assert [b for b, _ in prog.grad_final] == [0, 320, 346496]
    assert prog.grad_final[0][1] == len(prog.backward) and prog.grad_final[1][1] < len(prog.backward)
    # nothing scattered: everything after the first layer is ONE bucket, the first layer + metric sums close
    assert early == [(prog.grad_final[1][1], 320, lay.n_params, False)]
    assert late == [(0, 320), (lay.n_params, lay.n_total)]
    _covers_everything_once(lay, early, late)
    # with the dense-64 kernel reduce-scattered by its GEMM, the small remainder next to the metric sums joins
    # the closing exchange instead of paying for its own launch
    lay, prog, (early, late) = _plan(tf.models.mnist_cnn(), (28, 28, 1), scattered=[(320, 320 + 5408 * 64)])
    assert [(b, e, sc) for _, b, e, sc in early] == [(320, 346432, True)]
    assert late == [(0, 320), (346432, lay.n_total)]
    _covers_everything_once(lay, early, late)


def test_resnet50_buckets_run_from_the_last_layer_to_the_first(monkeypatch):
    model = tf.models.resnet50()
    lay, prog, (early, late) = _plan(model, (224, 224, 3))
    assert len(early) >= 4 and all(not sc for *_, sc in early)
    assert [n for n, *_ in early] == sorted(n for n, *_ in early)                 # issue order = finality order
    assert all((e - b) * 4 >= 16 << 20 for _, b, e, _ in early[:-1])              # 16 MB except the remainder
    assert late == [(0, 9408), (lay.n_params, lay.n_total)]                       # 7x7x3x64 stem + metric sums
    _covers_everything_once(lay, early, late)
    lay, prog, (finer, _) = _plan(model, (224, 224, 3), bucket_mb=4, monkeypatch=monkeypatch)
    assert len(finer) > len(early)


def test_single_parametrised_layer_has_only_a_closing_bucket():
    L = tf.keras.layers
    m = tf.keras.Sequential([L.Flatten(input_shape=(4, 4, 1)), L.Dense(10)])
    lay, prog, (early, late) = _plan(m, (4, 4, 1))
    assert early == [] and late == [(0, lay.n_params), (lay.n_params, lay.n_total)]
