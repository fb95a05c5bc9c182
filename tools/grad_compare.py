"""One SGD step with lr=1 on CPU and CUDA backends from the same init: weight deltas = gradients."""
import sys
import numpy as np
sys.path.insert(0, ".")
import distributed_b200 as tf
from distributed_b200.parallel.strategy import OneDeviceStrategy

rng = np.random.default_rng(0)
x = rng.random((64, 28, 28, 1)).astype(np.float32)
y = rng.integers(0, 10, 64)
res = {}
for be in ("cpu", "cuda"):
    with OneDeviceStrategy(backend=be).scope():
        m = tf.models.mnist_cnn(learning_rate=1.0)
    w0 = m.get_weights()
    m.fit(x, y, batch_size=64, epochs=1, steps_per_epoch=1, verbose=0)
    res[be] = [a - b for a, b in zip(w0, m.get_weights())]
for name, gc, gg in zip(m.weight_names(), res["cpu"], res["cuda"]):
    print(f"{name:18s} |g|max={np.abs(gc).max():.3e}  max|diff|={np.abs(gc - gg).max():.3e}  rel={np.abs(gc - gg).max() / np.abs(gc).max():.3e}")
