"""Metrics: only 'accuracy' exists in the reference (README.md:302).

Accuracy is a streaming count / samples-seen that resets every epoch (README.md:309-311) and,
under a strategy, is aggregated over all replicas (README.md:413: 98/1280).  The counts ride
in the tail slots of the gradient arena, so they cost no extra collective.
"""
from __future__ import annotations


def normalize(metrics) -> list[str]:
    if metrics is None:
        return []
    if isinstance(metrics, str):
        metrics = [metrics]
    out = []
    for m in metrics:
        if m in ("accuracy", "acc", "sparse_categorical_accuracy"):
            out.append("accuracy")
        else:
            raise ValueError(f"unknown metric {m!r}; the reference uses 'accuracy' only")
    return out
