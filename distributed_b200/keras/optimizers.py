"""Optimizers (README.md:302: ``SGD(learning_rate=0.001)``).

The update itself never runs as a separate kernel when more than one replica trains: it is
fused into the gradient all-reduce (average by 1/num_workers, then ``w -= lr * g``).
"""
from __future__ import annotations


class Optimizer:
    name = "optimizer"


class SGD(Optimizer):
    name = "SGD"

    def __init__(self, learning_rate: float = 0.01, momentum: float = 0.0, nesterov: bool = False,
                 name: str = "SGD", lr: float | None = None):
        if lr is not None:
            learning_rate = lr
        if momentum != 0.0 or nesterov:
            raise NotImplementedError("plain SGD only (momentum=0), as in the reference")
        self.learning_rate = float(learning_rate)
        self.momentum = 0.0
        self.name = name

    def get_config(self):
        return {"name": self.name, "learning_rate": self.learning_rate, "momentum": 0.0}


def get(identifier) -> Optimizer:
    if isinstance(identifier, Optimizer):
        return identifier
    if isinstance(identifier, str) and identifier.lower() == "sgd":
        return SGD()
    raise ValueError(f"unknown optimizer {identifier!r}")
