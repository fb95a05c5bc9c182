"""Flat parameter / gradient arenas.

All trainable tensors of a model live in ONE flat fp32 buffer in Keras layer order (so export
is a slice walk), shadowed by a same-shaped gradient buffer.
Four tail slots after the parameters carry the step's metric sums — loss sum, correct count,
sample count, spare — so that the loss/accuracy aggregation across replicas rides in the
gradient all-reduce instead of costing the extra collectives TensorFlow issues
(README.md:404-407).  In the weight arena the same tail slots hold the running epoch sums.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

ALIGN = 32          # floats (128 B): every tensor starts on a cache-line / TMA-friendly boundary
TAIL = 4            # loss_sum, correct, samples, spare


@dataclass
class Slot:
    name: str
    shape: tuple[int, ...]
    offset: int
    size: int
    layer_index: int


class ArenaLayout:
    def __init__(self, model_layers):
        from ..keras.layers import flatten_layers
        self.slots: list[Slot] = []
        self.leaves = flatten_layers(model_layers)
        self._index = {id(l): i for i, l in enumerate(self.leaves)}
        off = 0
        for li, layer in enumerate(self.leaves):
            for spec in layer.param_specs():
                self.slots.append(Slot(spec.name, tuple(spec.shape), off, spec.size, li))
                off += (spec.size + ALIGN - 1) // ALIGN * ALIGN
        self.n_params = off                 # padded parameter span (padding stays zero)
        self.n_total = off + TAIL
        self.count = sum(s.size for s in self.slots)   # true parameter count (347,146 for the README model)

    def slot(self, name: str) -> Slot:
        for s in self.slots:
            if s.name == name:
                return s
        raise KeyError(name)

    def view(self, flat, slot: Slot):
        """View of one tensor inside a flat torch buffer."""
        return flat[slot.offset:slot.offset + slot.size].view(slot.shape)

    def layer_slots(self, layer) -> list[Slot]:
        """Slots of a leaf layer (by object, or by its index in the flattened layer list)."""
        li = layer if isinstance(layer, int) else self._index[id(layer)]
        return [s for s in self.slots if s.layer_index == li]

    def leaf_index(self, layer) -> int:
        return self._index[id(layer)]

    def initial_values(self, model_layers, seed: int) -> np.ndarray:
        """Identical on every replica: one generator seeded per tensor from (seed, position)."""
        from ..keras.layers import initialize
        flat = np.zeros(self.n_total, np.float32)
        k = 0
        for layer in self.leaves:
            for spec in layer.param_specs():
                s = self.slots[k]
                rng = np.random.default_rng([seed, k])
                flat[s.offset:s.offset + s.size] = initialize(spec, rng).reshape(-1)
                k += 1
        return flat
