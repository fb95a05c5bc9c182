"""Residual networks through the same Sequential / fit API: one SGD step of a small ResNet
(projection + identity bottlenecks, strided 'same' convolutions, BatchNorm+ReLU, 3x3/2 max-pool,
global average pool) must match the same network written in PyTorch autograd."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


from _torch_ref import torch_forward as _torch_forward


@pytest.mark.parametrize("conv_path", ["tc", "direct"])
def test_small_resnet_step_matches_pytorch(conv_path, monkeypatch):
    import distributed_b200 as tf
    from distributed_b200.keras import layers as L
    if conv_path == "direct":
        monkeypatch.setenv("B200DIST_CONV", "direct")
    monkeypatch.setenv("B200DIST_DENSE", "simt")           # exact fp32 dense layers: compare tightly
    lr = 0.05
    layers = [L.Conv2D(32, 7, strides=2, padding="same", use_bias=False, input_shape=(32, 32, 3)),
              L.BatchNormalization(), L.ReLU(), L.MaxPooling2D(3, strides=2, padding="same"),
              tf.models.bottleneck(32, 1, project=True), tf.models.bottleneck(32, 1, project=False),
              tf.models.bottleneck(64, 2, project=True),
              L.GlobalAveragePooling2D(), L.Dense(10)]
    model = tf.keras.Sequential(layers)
    model.compile(loss=tf.keras.losses.SparseCategoricalCrossentropy(from_logits=True),
                  optimizer=tf.keras.optimizers.SGD(lr), metrics=["accuracy"])
    rng = np.random.default_rng(0)
    B = 16
    x = rng.standard_normal((B, 32, 32, 3)).astype(np.float32)
    y = rng.integers(0, 10, B)
    names = model.weight_names()
    w0 = {n: torch.tensor(a, device="cuda", requires_grad=not n.split("/")[-1].startswith("moving"))
          for n, a in zip(names, model.get_weights())}
    hist = model.fit(x, y, batch_size=B, epochs=1, steps_per_epoch=1, verbose=0)
    plan = " ".join(model._trainer.describe())
    assert ("tcgen05" in plan) == (conv_path == "tc"), plan
    running = {}
    logits = _torch_forward(w0, torch.tensor(x, device="cuda").permute(0, 3, 1, 2), model, running=running)
    loss = F.cross_entropy(logits, torch.tensor(y, device="cuda"))
    loss.backward()
    assert abs(hist.history["loss"][0] - loss.item()) < (3e-3 if conv_path == "tc" else 1e-4)
    w1 = dict(zip(names, model.get_weights()))
    # tf32 convolutions vs exact fp32: BatchNorm gradients sum to zero per channel, so the filter
    # gradients are differences of large terms and amplify the tf32 rounding of the operands
    for n in names:
        if n in running:            # moving statistics: updated through the optimizer (pseudo-gradient)
            assert np.allclose(w1[n], running[n].cpu().numpy(), atol=2e-3 if conv_path == "tc" else 1e-5, rtol=1e-3), n
            continue
        step = (lr * w0[n].grad).detach().cpu().numpy()
        ref = w0[n].detach().cpu().numpy() - step
        if conv_path == "tc":       # aggregate (L2) agreement of the update
            rel = np.linalg.norm(w1[n] - ref) / max(np.linalg.norm(step), 1e-12)
            assert rel < 0.2, (n, rel)
        else:                        # exact fp32 path: element-wise
            assert np.abs(w1[n] - ref).max() <= 2e-3 * np.abs(step).max() + 1e-6, n


def test_resnet50_builds_and_steps():
    import distributed_b200 as tf
    m = tf.models.resnet50(input_shape=(64, 64, 3), classes=100, learning_rate=0.01)
    assert m.count_params() > 23_000_000
    rng = np.random.default_rng(1)
    x = rng.standard_normal((16, 64, 64, 3)).astype(np.float32)
    y = rng.integers(0, 100, 16)
    h = m.fit(x, y, batch_size=8, epochs=2, steps_per_epoch=1, verbose=0)
    assert np.isfinite(h.history["loss"][0]) and 3.0 < h.history["loss"][0] < 8.0      # ~ ln(100) at init
    assert np.isfinite(h.history["loss"][1])
    plan = " ".join(m._trainer.describe())
    assert "tcgen05 tf32 implicit GEMM" in plan and "batchnorm+relu" in plan and "residual block" in plan


def test_small_channel_conv_runs_on_the_gemm_core_via_im2col():
    """A 3-channel 7x7/2 stem: patch matrix + tensor-core GEMMs must match the direct kernels."""
    from distributed_b200.engine.cuda_backend import CudaBackend
    g = torch.Generator(device="cuda").manual_seed(4)
    x = torch.randn((8, 64, 64, 3), device="cuda", generator=g)
    w = torch.randn((7, 7, 3, 64), device="cuda", generator=g) * 0.1
    dy = torch.randn((8, 32, 32, 64), device="cuda", generator=g)
    outs = {}
    for mode in ("direct", "im2col"):
        import os
        os.environ["B200DIST_CONV"] = "direct" if mode == "direct" else ""
        be = CudaBackend("cuda:0")
        y = be.alloc((8, 32, 32, 64))
        dw = be.alloc((7, 7, 3, 64))
        be.conv2d_fwd(x, w, None, y, (2, 2), (2, 2), True)()
        be.conv2d_wgrad(x, dy, dw, None, (2, 2), (2, 2))()
        torch.cuda.synchronize()
        outs[mode] = (y, dw)
    os.environ["B200DIST_CONV"] = ""
    assert (outs["direct"][0] - outs["im2col"][0]).abs().max().item() < 3e-3 * outs["direct"][0].abs().max().item()
    assert (outs["direct"][1] - outs["im2col"][1]).abs().max().item() < 3e-3 * outs["direct"][1].abs().max().item()


def test_small_resnet_step_bf16_policy_tracks_fp32(monkeypatch):
    """mixed_bfloat16: same network, same data, same initial weights as the fp32 policy; the loss and
    the parameter update must agree to bf16 accuracy (aggregate L2 per tensor), the moving statistics
    closely, and the plan must actually be the bf16 tensor-core path."""
    import distributed_b200 as tf
    from distributed_b200.keras import layers as L
    monkeypatch.setenv("B200DIST_DENSE", "simt")
    lr = 0.05
    rng = np.random.default_rng(0)
    B = 16
    x = rng.standard_normal((B, 32, 32, 3)).astype(np.float32)
    y = rng.integers(0, 10, B)
    out = {}
    for policy in ("float32", "mixed_bfloat16"):
        tf.keras.mixed_precision.set_global_policy(policy)
        try:
            layers = [L.Conv2D(64, 7, strides=2, padding="same", use_bias=False, input_shape=(32, 32, 3)),
                      L.BatchNormalization(), L.ReLU(), L.MaxPooling2D(3, strides=2, padding="same"),
                      tf.models.bottleneck(64, 1, project=True), tf.models.bottleneck(64, 1, project=False),
                      tf.models.bottleneck(128, 2, project=True),
                      L.GlobalAveragePooling2D(), L.Dense(10)]
            model = tf.keras.Sequential(layers)
            model.compile(loss=tf.keras.losses.SparseCategoricalCrossentropy(from_logits=True),
                          optimizer=tf.keras.optimizers.SGD(lr), metrics=["accuracy"])
            w_init = model.get_weights()
            hist = model.fit(x, y, batch_size=B, epochs=1, steps_per_epoch=1, verbose=0)
            out[policy] = (hist.history["loss"][0], w_init, model.get_weights(), " ".join(model._trainer.describe()),
                           model.weight_names())
        finally:
            tf.keras.mixed_precision.set_global_policy("float32")
    l32, wi32, w32, plan32, names = out["float32"]
    l16, wi16, w16, plan16, _ = out["mixed_bfloat16"]
    assert "tcgen05 bf16 implicit GEMM" in plan16 and "im2col + tcgen05 bf16" in plan16, plan16
    assert "bf16" not in plan32
    assert all(np.array_equal(a, b) for a, b in zip(wi32, wi16))
    assert abs(l16 - l32) < 1e-2, (l16, l32)
    # Calibrated with a bf16-rounding emulation of this network in PyTorch (tests/_torch_ref.py,
    # emulate_bf16=True): per-tensor update error vs fp32 is 0.17-0.35 (relative L2) with cosine >= 0.93 —
    # BatchNorm makes every gradient a difference of large terms.  A wrong kernel gives cosine ~ 0.
    rels, bad = [], {}
    for n, a0, a32, a16 in zip(names, wi32, w32, w16):
        if n.split("/")[-1].startswith("moving"):
            assert np.allclose(a16, a32, atol=2e-2, rtol=2e-2), n
            continue
        s32, s16 = (a32 - a0).ravel().astype(np.float64), (a16 - a0).ravel().astype(np.float64)
        rel = np.linalg.norm(s16 - s32) / max(np.linalg.norm(s32), 1e-12)
        cos = float(s16 @ s32) / max(np.linalg.norm(s16) * np.linalg.norm(s32), 1e-30)
        rels.append(rel)
        if rel > 0.7 or cos < 0.8:
            bad[n] = (round(rel, 3), round(cos, 3))
    assert not bad, bad
    assert np.median(rels) < 0.4, np.median(rels)


def test_resnet50_bf16_builds_and_steps():
    import distributed_b200 as tf
    tf.keras.mixed_precision.set_global_policy("mixed_bfloat16")
    try:
        m = tf.models.resnet50(input_shape=(64, 64, 3), classes=100, learning_rate=0.01)
        rng = np.random.default_rng(1)
        x = rng.standard_normal((16, 64, 64, 3)).astype(np.float32)
        y = rng.integers(0, 100, 16)
        h = m.fit(x, y, batch_size=8, epochs=2, steps_per_epoch=1, verbose=0)
        assert np.isfinite(h.history["loss"][0]) and 3.0 < h.history["loss"][0] < 8.0
        assert np.isfinite(h.history["loss"][1])
        plan = " ".join(m._trainer.describe())
        assert "tcgen05 bf16 implicit GEMM" in plan and "direct conv" not in plan
        p = m.predict(x[:8], batch_size=8)
        assert np.isfinite(p).all()
    finally:
        tf.keras.mixed_precision.set_global_policy("float32")
