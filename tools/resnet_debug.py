import os, sys
import numpy as np, torch, torch.nn.functional as F
sys.path.insert(0, ".")
sys.path.insert(0, "tests")
os.environ["B200DIST_DENSE"] = "simt"
import distributed_b200 as tf
from distributed_b200.keras import layers as L
from test_resnet_gpu import _torch_forward

rng = np.random.default_rng(0)
B = 16
x = rng.standard_normal((B, 32, 32, 3)).astype(np.float32)
y = rng.integers(0, 10, B)


def check(name, mk):
    for path in ("direct", "tc"):
        os.environ["B200DIST_CONV"] = "direct" if path == "direct" else ""
        model = tf.keras.Sequential(mk())
        model.compile(loss=tf.keras.losses.SparseCategoricalCrossentropy(from_logits=True),
                      optimizer=tf.keras.optimizers.SGD(0.05), metrics=["accuracy"])
        names = model.weight_names()
        w0 = {n: torch.tensor(a, device="cuda", requires_grad=True) for n, a in zip(names, model.get_weights())}
        h = model.fit(x, y, batch_size=B, epochs=1, steps_per_epoch=1, verbose=0)
        logits = _torch_forward(w0, torch.tensor(x, device="cuda").permute(0, 3, 1, 2), model)
        loss = F.cross_entropy(logits, torch.tensor(y, device="cuda"))
        loss.backward()
        w1 = dict(zip(names, model.get_weights()))
        worst = max(((np.abs(w1[n] - (w0[n] - 0.05 * w0[n].grad).detach().cpu().numpy()).max() /
                      (np.abs(0.05 * w0[n].grad.detach().cpu().numpy()).max() + 1e-12), n) for n in names))
        print(f"{name:34s} {path:6s} loss ours={h.history['loss'][0]:.5f} torch={loss.item():.5f}  worst weight rel err={worst[0]:.3e} ({worst[1]})")


stem = lambda: [L.Conv2D(32, 7, strides=2, padding="same", use_bias=False, input_shape=(32, 32, 3)), L.BatchNormalization(), L.ReLU()]
tail = lambda: [L.GlobalAveragePooling2D(), L.Dense(10)]
check("stem", lambda: stem() + tail())
check("stem+pool", lambda: stem() + [L.MaxPooling2D(3, strides=2, padding="same")] + tail())
check("stem+conv3x3same", lambda: stem() + [L.Conv2D(32, 3, padding="same", use_bias=False), L.BatchNormalization(), L.ReLU()] + tail())
check("stem+conv1x1", lambda: stem() + [L.Conv2D(64, 1, use_bias=False), L.BatchNormalization(), L.ReLU()] + tail())
check("stem+conv3x3/s2", lambda: stem() + [L.Conv2D(32, 3, strides=2, padding="same", use_bias=False), L.BatchNormalization(), L.ReLU()] + tail())
check("stem+bottleneck(project)", lambda: stem() + [tf.models.bottleneck(32, 1, True)] + tail())
check("stem+bottleneck(project)+identity", lambda: stem() + [tf.models.bottleneck(32, 1, True), tf.models.bottleneck(32, 1, False)] + tail())
check("stem+bottleneck(stride2)", lambda: stem() + [tf.models.bottleneck(32, 2, True)] + tail())
