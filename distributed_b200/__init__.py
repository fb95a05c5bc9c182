"""distributed_b200 — Blackwell-native multi-worker synchronous data-parallel training.

The public surface mirrors what Mrhs121/distributed's README uses from TensorFlow 2.0, so
``import distributed_b200 as tf`` runs the README's Python snippets unchanged
(README.md:281-305, 318-393).  Sub-modules are imported lazily to keep ``import`` cheap.
"""
from __future__ import annotations

import importlib

__version__ = "2.0.0-b200"
__all__ = ["keras", "distribute", "__version__"]

_LAZY = {
    "keras": "distributed_b200.keras",
    "distribute": "distributed_b200.parallel.distribute",
    "launch": "distributed_b200.launch",
    "ops": "distributed_b200.ops",
    "models": "distributed_b200.models",
    "utils": "distributed_b200.utils",
    "parallel": "distributed_b200.parallel",
    "engine": "distributed_b200.engine",
}


def __getattr__(name):
    if name in _LAZY:
        mod = importlib.import_module(_LAZY[name])
        globals()[name] = mod
        return mod
    raise AttributeError(f"module 'distributed_b200' has no attribute {name!r}")
