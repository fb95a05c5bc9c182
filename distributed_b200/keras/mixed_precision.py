"""``tf.keras.mixed_precision``: the dtype policy of the convolutional part of a model.

The reference trains in float32 (README.md:292-302 sets no policy); BASELINE.json config 4
names ResNet-50 in bf16, which is what ``set_global_policy("mixed_bfloat16")`` selects on the
CUDA backend:

* 4-D activations and their gradients are stored in bfloat16; convolutions run on the tcgen05
  tensor cores with bf16 operands and fp32 (TMEM) accumulation;
* variables stay float32 (``variable_dtype``), exactly like Keras' mixed policies: the optimizer —
  fused into the gradient all-reduce — updates fp32 master weights with fp32 gradients, and a
  bf16 copy of the weights is refreshed by one streaming kernel at the start of every step;
* BatchNormalization statistics, the loss and everything after Flatten /
  GlobalAveragePooling2D (the Dense head) are float32 / tf32.

The policy is read when a model's trainer is built (first ``fit`` / ``evaluate`` / ``predict``).
``B200DIST_DTYPE=bf16|fp32`` overrides it from the environment.  The CPU backend always computes
in float32.
"""
from __future__ import annotations

import os


class Policy:
    def __init__(self, name: str):
        if name not in ("float32", "mixed_bfloat16"):
            raise ValueError(f"unsupported dtype policy {name!r}: use 'float32' or 'mixed_bfloat16'")
        self.name = name

    @property
    def compute_dtype(self) -> str:
        return "bfloat16" if self.name == "mixed_bfloat16" else "float32"

    @property
    def variable_dtype(self) -> str:
        return "float32"

    def __repr__(self):
        return f'<Policy "{self.name}">'


_global = Policy("float32")


def set_global_policy(policy) -> None:
    global _global
    _global = policy if isinstance(policy, Policy) else Policy(str(policy))


def global_policy() -> Policy:
    env = os.environ.get("B200DIST_DTYPE", "").lower()
    if env in ("bf16", "bfloat16", "mixed_bfloat16"):
        return Policy("mixed_bfloat16")
    if env in ("fp32", "float32", "tf32"):
        return Policy("float32")
    return _global
