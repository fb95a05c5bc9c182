"""Losses (README.md:301: ``SparseCategoricalCrossentropy(from_logits=True)``)."""
from __future__ import annotations


class Loss:
    name = "loss"


class SparseCategoricalCrossentropy(Loss):
    """Integer labels vs logits; the engine fuses softmax, loss, gradient and accuracy count."""
    name = "sparse_categorical_crossentropy"

    def __init__(self, from_logits: bool = False, reduction: str = "auto", name: str | None = None):
        if not from_logits:
            raise NotImplementedError(
                "only from_logits=True is supported (the reference model ends in a linear Dense(10))")
        self.from_logits = True
        self.reduction = reduction
        if name:
            self.name = name


def get(identifier) -> Loss:
    if isinstance(identifier, Loss):
        return identifier
    if identifier in ("sparse_categorical_crossentropy",):
        raise ValueError("pass SparseCategoricalCrossentropy(from_logits=True): the string form "
                         "implies probabilities, which this framework does not take")
    raise ValueError(f"unknown loss {identifier!r}")
