from __future__ import annotations

from ..keras import layers as L
from ..keras.losses import SparseCategoricalCrossentropy
from ..keras.models import Sequential
from ..keras.optimizers import SGD


def mnist_cnn(compile: bool = True, learning_rate: float = 0.001) -> Sequential:
    """The README model, verbatim (README.md:292-302): 347,146 parameters in 6 tensors."""
    model = Sequential([
        L.Conv2D(32, 3, activation="relu", input_shape=(28, 28, 1)),
        L.MaxPooling2D(),
        L.Flatten(),
        L.Dense(64, activation="relu"),
        L.Dense(10),
    ])
    if compile:
        model.compile(loss=SparseCategoricalCrossentropy(from_logits=True),
                      optimizer=SGD(learning_rate=learning_rate), metrics=["accuracy"])
    return model


def scaled_cnn(compile: bool = True, learning_rate: float = 0.001) -> Sequential:
    """BASELINE.json config 3: the same CNN scaled to conv64 / conv128 / dense1024."""
    model = Sequential([
        L.Conv2D(64, 3, activation="relu", input_shape=(28, 28, 1)),
        L.MaxPooling2D(),
        L.Conv2D(128, 3, activation="relu"),
        L.MaxPooling2D(),
        L.Flatten(),
        L.Dense(1024, activation="relu"),
        L.Dense(10),
    ], name="scaled_cnn")
    if compile:
        model.compile(loss=SparseCategoricalCrossentropy(from_logits=True),
                      optimizer=SGD(learning_rate=learning_rate), metrics=["accuracy"])
    return model
