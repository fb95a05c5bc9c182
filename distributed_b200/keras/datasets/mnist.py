"""MNIST loader with the Keras signature (README.md:286-287; R ``dataset_mnist()`` README.md:51).

Keras downloads ``mnist.npz``; this environment has no network, so the loader looks for a
local copy (``$B200DIST_DATA``, ``~/.keras/datasets/mnist.npz``) and otherwise generates a
deterministic synthetic set of exactly MNIST's shape and dtype: 60000/10000 uint8 images of
28x28 with integer labels 0..9.  The synthetic digits are class-dependent stroke patterns
plus noise, so a model can actually learn them (useful for convergence tests).
"""
from __future__ import annotations

import os
from pathlib import Path

import numpy as np

_CACHE: dict = {}


def _find_local(path: str) -> Path | None:
    cands = []
    if os.environ.get("B200DIST_DATA"):
        cands.append(Path(os.environ["B200DIST_DATA"]) / path)
    cands.append(Path.home() / ".keras" / "datasets" / path)
    for c in cands:
        if c.exists():
            return c
    return None


def synthetic(n_train: int = 60000, n_test: int = 10000, seed: int = 0):
    """MNIST-shaped synthetic data: ((x_train, y_train), (x_test, y_test)), uint8 / uint8."""
    rng = np.random.default_rng(seed)
    # one 28x28 prototype per class: a few random strokes, fixed by the seed
    protos = np.zeros((10, 28, 28), np.float32)
    for c in range(10):
        prng = np.random.default_rng(1000 + c)
        for _ in range(4):
            r0, c0 = prng.integers(4, 24, 2)
            dr, dc = prng.integers(-1, 2, 2)
            if dr == 0 and dc == 0:
                dc = 1
            for t in range(prng.integers(6, 14)):
                r, cc = r0 + dr * t, c0 + dc * t
                if 1 <= r < 27 and 1 <= cc < 27:
                    protos[c, r - 1:r + 2, cc - 1:cc + 2] += 0.6
    protos = np.clip(protos, 0, 1)

    def make(n):
        y = rng.integers(0, 10, n).astype(np.uint8)
        x = protos[y] * 255.0
        x = x + rng.normal(0, 20.0, x.shape)
        return np.clip(x, 0, 255).astype(np.uint8), y

    return make(n_train), make(n_test)


def load_data(path: str = "mnist.npz"):
    key = ("mnist", path)
    if key not in _CACHE:
        local = _find_local(path)
        if local is not None:
            with np.load(local, allow_pickle=False) as f:
                _CACHE[key] = ((f["x_train"], f["y_train"]), (f["x_test"], f["y_test"]))
        else:
            _CACHE[key] = synthetic()
    return _CACHE[key]
