"""BASELINE.json config 1 on this host: the README CNN on the C++ CPU backend at world_size 1 and 2
(POSIX shared-memory all-reduce), plus the same model in stock PyTorch CPU eager for comparison.

    python tools/cpu_config_bench.py [steps]

Wall-clock per step after a warm-up fit (the container is shared: run it a few times and read the best)."""
import json
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))


def work(steps, ctx):
    import os
    os.environ["B200DIST_BACKEND"] = "cpu"
    os.environ["TF_CONFIG"] = json.dumps(ctx.tf_config(base_port=int(ctx.address[0].split(":")[1]) % 20000 + 20000))
    import numpy as np
    import distributed_b200 as tf
    strategy = tf.distribute.experimental.MultiWorkerMirroredStrategy()
    n = len(ctx.address)
    rng = np.random.default_rng(0)
    x = rng.random((64 * n * (steps + 10), 28, 28, 1)).astype(np.float32)
    y = rng.integers(0, 10, len(x))
    with strategy.scope():
        m = tf.models.mnist_cnn(learning_rate=0.001)
    m.fit(x, y, batch_size=64 * n, epochs=1, steps_per_epoch=10, verbose=0)
    t0 = time.perf_counter()
    m.fit(x, y, batch_size=64 * n, epochs=1, steps_per_epoch=steps, verbose=0)
    t = time.perf_counter() - t0
    return f"{1e3 * t / steps:.2f} ms/step, {64 * n * steps / t:.0f} images/s (global batch {64 * n})"


def torch_cpu(steps):
    import torch
    import torch.nn as nn
    import torch.nn.functional as F
    net = nn.Sequential(nn.Conv2d(1, 32, 3), nn.ReLU(), nn.MaxPool2d(2), nn.Flatten(), nn.Linear(5408, 64), nn.ReLU(),
                        nn.Linear(64, 10))
    opt = torch.optim.SGD(net.parameters(), lr=1e-3)
    x, y = torch.rand(64 * (steps + 3), 1, 28, 28), torch.randint(0, 10, (64 * (steps + 3),))

    def step(i):
        opt.zero_grad()
        F.cross_entropy(net(x[i * 64:(i + 1) * 64]), y[i * 64:(i + 1) * 64]).backward()
        opt.step()
    for i in range(3):
        step(i)
    t0 = time.perf_counter()
    for i in range(3, steps + 3):
        step(i)
    t = time.perf_counter() - t0
    return f"{1e3 * t / steps:.2f} ms/step, {64 * steps / t:.0f} images/s ({torch.get_num_threads()} threads)"


if __name__ == "__main__":
    from distributed_b200.launch import barrier_apply
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 50
    for world in (1, 2):
        print(f"distributed_b200 CPU backend, world_size {world}:", barrier_apply(work, world, args=[steps] * world, timeout=600)[0])
    print("stock PyTorch CPU eager, 1 process:", torch_cpu(steps))
