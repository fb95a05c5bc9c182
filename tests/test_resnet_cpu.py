"""Residual networks on the C++ CPU backend: the lowering of BatchNormalization (+ReLU, moving
statistics as optimizer-updated slots), Residual blocks (shared output gradient, masked gradient
sum), strided 'same' convolutions, general max pooling and global average pooling must match the
same network in PyTorch autograd — the CPU tier of SURVEY.md section 4.2 for BASELINE config 4."""
import numpy as np
import torch
import torch.nn.functional as F

from _torch_ref import torch_forward


def _small_resnet(tf, L, lr):
    layers = [L.Conv2D(8, 3, strides=2, padding="same", use_bias=False, input_shape=(16, 16, 3)),
              L.BatchNormalization(momentum=0.9), L.ReLU(), L.MaxPooling2D(3, strides=2, padding="same"),
              tf.models.bottleneck(8, 1, project=True), tf.models.bottleneck(8, 1, project=False),
              tf.models.bottleneck(16, 2, project=True),
              L.GlobalAveragePooling2D(), L.Dense(10)]
    model = tf.keras.Sequential(layers)
    model.compile(loss=tf.keras.losses.SparseCategoricalCrossentropy(from_logits=True),
                  optimizer=tf.keras.optimizers.SGD(lr), metrics=["accuracy"])
    return model


def test_small_resnet_step_on_cpu_matches_pytorch():
    import distributed_b200 as tf
    from distributed_b200.keras import layers as L
    lr = 0.05
    model = _small_resnet(tf, L, lr)
    rng = np.random.default_rng(0)
    B = 8
    x = rng.standard_normal((B, 16, 16, 3)).astype(np.float32)
    y = rng.integers(0, 10, B)
    names = model.weight_names()
    assert any(n.endswith("moving_variance:0") for n in names)
    w0 = {n: torch.tensor(a, requires_grad=not n.split("/")[-1].startswith("moving"))
          for n, a in zip(names, model.get_weights())}
    hist = model.fit(x, y, batch_size=B, epochs=1, steps_per_epoch=1, verbose=0)
    plan = " ".join(model._trainer.describe())
    assert "residual block" in plan and "batchnorm+relu" in plan
    running = {}
    logits = torch_forward(w0, torch.tensor(x).permute(0, 3, 1, 2), model, running=running)
    loss = F.cross_entropy(logits, torch.tensor(y))
    loss.backward()
    assert abs(hist.history["loss"][0] - loss.item()) < 1e-4
    w1 = dict(zip(names, model.get_weights()))
    for n in names:
        if n in running:        # moving statistics: momentum * old + (1 - momentum) * batch statistic
            assert np.allclose(w1[n], running[n].numpy(), atol=1e-5, rtol=1e-4), n
            continue
        step = (lr * w0[n].grad).numpy()
        ref = w0[n].detach().numpy() - step
        assert np.abs(w1[n] - ref).max() <= 2e-3 * np.abs(step).max() + 1e-6, n


def test_batchnorm_inference_uses_moving_statistics_on_cpu():
    import distributed_b200 as tf
    from distributed_b200.keras import layers as L
    model = _small_resnet(tf, L, 0.05)
    rng = np.random.default_rng(1)
    B = 8
    x = rng.standard_normal((4 * B, 16, 16, 3)).astype(np.float32)
    y = rng.integers(0, 10, 4 * B)
    model.fit(x, y, batch_size=B, epochs=1, steps_per_epoch=4, verbose=0)      # moves the statistics
    names = model.weight_names()
    w = {n: torch.tensor(a) for n, a in zip(names, model.get_weights())}
    assert not np.allclose(w["batch_normalization/moving_mean:0"].numpy(), 0.0)
    pred = model.predict(x[:B], batch_size=B)
    ref = torch_forward(w, torch.tensor(x[:B]).permute(0, 3, 1, 2), model, training=False).numpy()
    assert np.allclose(pred, ref, atol=2e-4, rtol=1e-3)
    # a single-sample change must not move the other samples' outputs (no batch statistics at inference)
    x2 = x[:B].copy()
    x2[0] += 1.0
    pred2 = model.predict(x2, batch_size=B)
    assert np.allclose(pred2[1:], pred[1:], atol=1e-6)


def test_weights_roundtrip_keeps_moving_statistics(tmp_path):
    import distributed_b200 as tf
    from distributed_b200.keras import layers as L
    model = _small_resnet(tf, L, 0.05)
    rng = np.random.default_rng(2)
    x = rng.standard_normal((8, 16, 16, 3)).astype(np.float32)
    y = rng.integers(0, 10, 8)
    model.fit(x, y, batch_size=8, epochs=1, steps_per_epoch=1, verbose=0)
    path = str(tmp_path / "resnet.npz")
    model.save_weights(path)
    other = _small_resnet(tf, L, 0.05)
    other.build((None, 16, 16, 3))
    other.load_weights(path)
    for a, b, n in zip(model.get_weights(), other.get_weights(), model.weight_names()):
        assert np.array_equal(a, b), n
    assert np.allclose(model.predict(x, batch_size=8), other.predict(x, batch_size=8), atol=1e-6)


def test_resnet_builder_lowers_with_fused_block_tails_and_gradient_sums():
    """The ResNet builder (one bottleneck per stage) through fit() on the CPU backend: every block's tail is
    the BatchNorm+add+ReLU fusion, every block's gradient sum rides in its first convolution's data gradient,
    weights in Keras order include the moving statistics, and the loss starts near ln(classes)."""
    import distributed_b200 as tf
    m = tf.models.resnet(blocks=(1, 1, 1, 1), input_shape=(32, 32, 3), classes=10, width=8, learning_rate=0.01)
    names = m.weight_names()
    assert names[0] == "conv2d/kernel:0" and names[1:5] == [
        "batch_normalization/gamma:0", "batch_normalization/beta:0", "batch_normalization/moving_mean:0",
        "batch_normalization/moving_variance:0"]
    rng = np.random.default_rng(3)
    x = rng.standard_normal((8, 32, 32, 3)).astype(np.float32)
    y = rng.integers(0, 10, 8)
    h = m.fit(x, y, batch_size=4, epochs=1, steps_per_epoch=2, verbose=0)
    assert np.isfinite(h.history["loss"][0]) and 1.5 < h.history["loss"][0] < 4.0
    plan = m._trainer.describe()
    assert sum("residual block" in p for p in plan) == 4
    assert sum("inside the last batchnorm" in p for p in plan) == 4
    assert sum("gradient sum fused" in p for p in plan) == 4
    assert sum(p.startswith("batchnorm+residual+relu") for p in plan) == 4
    trainable = sum(l.count_trainable_params() for l in m.layers)
    assert 0 < m.count_params() - trainable == sum(
        int(np.prod(a.shape)) for n, a in zip(names, m.get_weights()) if "/moving_" in n)
