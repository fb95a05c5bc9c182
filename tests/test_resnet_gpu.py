"""Residual networks through the same Sequential / fit API: one SGD step of a small ResNet
(projection + identity bottlenecks, strided 'same' convolutions, BatchNorm+ReLU, 3x3/2 max-pool,
global average pool) must match the same network written in PyTorch autograd."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _torch_forward(w, x, model):
    """Replays the framework's layer list with torch ops; `w` maps weight name -> tensor."""
    from distributed_b200.keras import layers as L

    def conv(layer, t):
        k = w[layer.name + "/kernel:0"].permute(3, 2, 0, 1)
        b = w.get(layer.name + "/bias:0")
        ph, pw = layer.pad_amounts()
        # TF 'same' with stride 2 pads asymmetrically (extra pixel at the end)
        H, W = t.shape[2], t.shape[3]
        oh, ow = layer.output_shape[0], layer.output_shape[1]
        th = max((oh - 1) * layer.strides[0] + layer.kernel_size[0] - H, 0)
        tw = max((ow - 1) * layer.strides[1] + layer.kernel_size[1] - W, 0)
        t = F.pad(t, (pw, tw - pw, ph, th - ph))
        t = F.conv2d(t, k, b, stride=layer.strides)
        return torch.relu(t) if layer.activation == "relu" else t

    def run(layers, t):
        for layer in layers:
            if isinstance(layer, L.Conv2D):
                t = conv(layer, t)
            elif isinstance(layer, L.BatchNormalization):
                t = F.batch_norm(t, None, None, w[layer.name + "/gamma:0"], w[layer.name + "/beta:0"],
                                 training=True, eps=layer.epsilon)
                if layer.activation == "relu":
                    t = torch.relu(t)
            elif isinstance(layer, L.MaxPooling2D):
                p = layer.pad_amount()
                k, s = layer.pool_size[0], layer.strides[0]
                H = t.shape[2]
                total = max((layer.output_shape[0] - 1) * s + k - H, 0)
                t = F.pad(t, (p, total - p, p, total - p), value=float("-inf"))
                t = F.max_pool2d(t, k, s)
            elif isinstance(layer, L.Residual):
                m = run(layer.main, t)
                sc = run(layer.shortcut, t) if layer.shortcut else t
                t = torch.relu(m + sc)
            elif isinstance(layer, L.GlobalAveragePooling2D):
                t = t.mean((2, 3))
            elif isinstance(layer, L.Flatten):
                t = t.permute(0, 2, 3, 1).reshape(t.shape[0], -1)
            elif isinstance(layer, L.Dense):
                t = t @ w[layer.name + "/kernel:0"] + w[layer.name + "/bias:0"]
                if layer.activation == "relu":
                    t = torch.relu(t)
            else:
                raise AssertionError(type(layer))
        return t
    return run(model.layers, x)


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
    w0 = {n: torch.tensor(a, device="cuda", requires_grad=True) for n, a in zip(names, model.get_weights())}
    hist = model.fit(x, y, batch_size=B, epochs=1, steps_per_epoch=1, verbose=0)
    plan = " ".join(model._trainer.describe())
    assert ("tcgen05" in plan) == (conv_path == "tc"), plan
    logits = _torch_forward(w0, torch.tensor(x, device="cuda").permute(0, 3, 1, 2), model)
    loss = F.cross_entropy(logits, torch.tensor(y, device="cuda"))
    loss.backward()
    assert abs(hist.history["loss"][0] - loss.item()) < (3e-3 if conv_path == "tc" else 1e-4)
    w1 = dict(zip(names, model.get_weights()))
    # tf32 convolutions vs exact fp32: BatchNorm gradients sum to zero per channel, so the filter
    # gradients are differences of large terms and amplify the tf32 rounding of the operands
    for n in names:
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
