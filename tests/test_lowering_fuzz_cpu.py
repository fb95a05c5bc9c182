"""Random small architectures (convolutions with / without bias and ReLU, strides, 'same' / 'valid', pooling
variants, BatchNorm with / without ReLU, residual blocks with identity or projection shortcuts and with or
without the block ReLU, Flatten or global average pooling, one or two Dense layers) lowered by the engine on
the CPU backend: one SGD step must match PyTorch autograd on every weight, moving statistics included.
This pins the lowering rules (who masks which gradient, fused block tails, fused gradient sums, stage order)
independently of the CUDA kernels."""
import numpy as np
import torch
import torch.nn.functional as F
from hypothesis import given, settings, strategies as st

from _torch_ref import torch_forward


def _build(tf, L, draw, widths=(4, 8)):
    layers, c, hw = [], 3, 12
    first = True

    def conv(filters, k, stride=1, padding="same", relu=False, bias=True):
        nonlocal first
        kw = {"input_shape": (12, 12, 3)} if first else {}
        first = False
        return L.Conv2D(filters, k, strides=stride, padding=padding, use_bias=bias,
                        activation="relu" if relu else None, **kw)

    c = draw(st.sampled_from(list(widths)))
    layers.append(conv(c, draw(st.sampled_from([1, 3])), relu=draw(st.booleans()), bias=draw(st.booleans())))
    for _ in range(draw(st.integers(1, 3))):
        kind = draw(st.sampled_from(["conv", "conv_valid", "bn", "pool", "relu", "res_id", "res_proj"]))
        if kind == "conv":
            stride = draw(st.sampled_from([1, 2])) if hw >= 6 else 1
            c2 = draw(st.sampled_from(list(widths)))
            layers.append(conv(c2, draw(st.sampled_from([1, 3])), stride, "same", draw(st.booleans()), draw(st.booleans())))
            c, hw = c2, -(-hw // stride)
        elif kind == "conv_valid" and hw >= 5:
            c2 = draw(st.sampled_from(list(widths)))
            layers.append(conv(c2, 3, 1, "valid", draw(st.booleans()), draw(st.booleans())))
            c, hw = c2, hw - 2
        elif kind == "relu":
            layers.append(L.ReLU())          # folded into a producer without activation, else a stand-alone layer
        elif kind == "bn":
            layers.append(L.BatchNormalization(momentum=0.9))
            if draw(st.booleans()):
                layers.append(L.ReLU())
        elif kind == "pool" and hw >= 4:
            if draw(st.booleans()):
                layers.append(L.MaxPooling2D())
                hw //= 2
            else:
                layers.append(L.MaxPooling2D(3, strides=2, padding="same"))
                hw = -(-hw // 2)
        elif kind in ("res_id", "res_proj"):
            proj = kind == "res_proj"
            c2 = draw(st.sampled_from(list(widths))) if proj else c
            stride = draw(st.sampled_from([1, 2])) if (proj and hw >= 6) else 1
            main = [conv(4, 1, bias=False), L.BatchNormalization(momentum=0.9), L.ReLU(),
                    conv(c2, 3, stride, "same", bias=draw(st.booleans()))]
            if draw(st.booleans()):
                main.append(L.BatchNormalization(momentum=0.9))      # fused block tail when it ends the branch
            shortcut = [conv(c2, 1, stride, "same", bias=False), L.BatchNormalization(momentum=0.9)] if proj else None
            layers.append(L.Residual(main, shortcut, activation="relu" if draw(st.booleans()) else None))
            c, hw = c2, -(-hw // stride)
    layers.append(L.GlobalAveragePooling2D() if draw(st.booleans()) else L.Flatten())
    if draw(st.booleans()):
        layers.append(L.Dense(6, activation="relu"))
    layers.append(L.Dense(5))
    return layers


@settings(max_examples=40, deadline=None)
@given(data=st.data())
def test_random_architectures_match_autograd(data):
    import distributed_b200 as tf
    from distributed_b200.keras import layers as L
    lr = 0.05
    model = tf.keras.Sequential(_build(tf, L, data.draw))
    model.compile(loss=tf.keras.losses.SparseCategoricalCrossentropy(from_logits=True),
                  optimizer=tf.keras.optimizers.SGD(lr), metrics=["accuracy"])
    rng = np.random.default_rng(data.draw(st.integers(0, 1000)))
    B = 4
    x = rng.standard_normal((B, 12, 12, 3)).astype(np.float32)
    y = rng.integers(0, 5, B)
    import os
    if os.environ.get("B200DIST_FUZZ_LOG"):
        with open(os.environ["B200DIST_FUZZ_LOG"], "a") as f:
            f.write(repr([(type(l).__name__, getattr(l, "kernel_size", None), getattr(l, "strides", None),
                           getattr(l, "filters", None), getattr(l, "activation", None),
                           [(type(m).__name__, getattr(m, "filters", None), getattr(m, "strides", None)) for m in
                            getattr(l, "main", [])], len(getattr(l, "shortcut", []))) for l in model.layers]) + "\n")
    names = model.weight_names()
    w0 = {n: torch.tensor(a, requires_grad=not n.split("/")[-1].startswith("moving"))
          for n, a in zip(names, model.get_weights())}
    hist = model.fit(x, y, batch_size=B, epochs=1, steps_per_epoch=1, verbose=0)
    running = {}
    logits = torch_forward(w0, torch.tensor(x).permute(0, 3, 1, 2), model, running=running)
    loss = F.cross_entropy(logits, torch.tensor(y))
    loss.backward()
    plan = model._trainer.describe()
    assert abs(hist.history["loss"][0] - loss.item()) < 2e-4, plan
    w1 = dict(zip(names, model.get_weights()))
    for n in names:
        if n in running:
            assert np.allclose(w1[n], running[n].numpy(), atol=1e-5, rtol=1e-4), (n, plan)
            continue
        g = w0[n].grad
        step = (lr * g).numpy() if g is not None else np.zeros_like(w1[n])
        ref = w0[n].detach().numpy() - step
        assert np.abs(w1[n] - ref).max() <= 5e-3 * max(np.abs(step).max(), 1e-4) + 2e-6, (n, plan)
