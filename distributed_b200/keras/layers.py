"""Keras-style layer objects (README.md:292-298; R equivalents README.md:58-68).

Layers here are *descriptions*: hyper-parameters, shape inference and parameter specs in
Keras' own order and layouts (Conv2D kernel HWIO then bias; Dense kernel (in, out) then bias;
activations NHWC; Flatten in H, W, C order).  The numerical work happens in the engine, which
lowers a stack of layers to fused sm_100a kernels (or to the C++ CPU backend).
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np

_ACTIVATIONS = (None, "linear", "relu")


def _pair(v) -> tuple[int, int]:
    if isinstance(v, (tuple, list)):
        assert len(v) == 2
        return int(v[0]), int(v[1])
    return int(v), int(v)


def _norm_activation(a):
    if callable(a):
        a = getattr(a, "__name__", None)
    if a not in _ACTIVATIONS:
        raise ValueError(f"activation {a!r} is not supported (use None or 'relu')")
    return None if a == "linear" else a


@dataclass
class ParamSpec:
    name: str                      # Keras weight name, e.g. "conv2d/kernel:0"
    shape: tuple[int, ...]
    init: str                      # "glorot_uniform" | "zeros" | "ones"
    fan_in: int = 0
    fan_out: int = 0
    trainable: bool = True

    @property
    def size(self) -> int:
        return int(np.prod(self.shape))


class Layer:
    """Base class; subclasses define ``compute_output_shape`` and ``param_specs``."""

    _counters: dict[str, int] = {}

    def __init__(self, name: str | None = None, input_shape=None, **kwargs):
        if kwargs:
            raise TypeError(f"unsupported arguments for {type(self).__name__}: {sorted(kwargs)}")
        self._given_name = name
        self.name = name  # finalised by the model (conv2d, conv2d_1, ... per model)
        self.input_shape_arg = tuple(input_shape) if input_shape is not None else None
        self.input_shape: tuple[int, ...] | None = None    # without batch dimension
        self.output_shape: tuple[int, ...] | None = None
        self.built = False

    base_name = "layer"
    activation = None

    def build(self, input_shape: tuple[int, ...]) -> tuple[int, ...]:
        self.input_shape = tuple(int(d) for d in input_shape)
        self.output_shape = self.compute_output_shape(self.input_shape)
        self.built = True
        return self.output_shape

    def compute_output_shape(self, input_shape):
        return input_shape

    def param_specs(self) -> list[ParamSpec]:
        return []

    def count_params(self) -> int:
        return sum(p.size for p in self.param_specs())

    def count_trainable_params(self) -> int:
        return sum(p.size for p in self.param_specs() if p.trainable)

    # -- serialisation (Keras-2 "config" dictionaries; model.save / load_model, README.md:237-246) ----------
    def _base_config(self) -> dict:
        cfg = {"name": self.name, "trainable": True, "dtype": "float32"}
        if self.input_shape_arg is not None:
            cfg["batch_input_shape"] = [None, *self.input_shape_arg]
        return cfg

    def get_config(self) -> dict:
        return self._base_config()

    @classmethod
    def from_config(cls, config: dict):
        """Rebuild a layer from ``get_config()`` output.  Keys of a genuine Keras config that this framework
        has no use for (dtype, trainable, data_format, regularizers, ...) are accepted and dropped; values it
        cannot honour (e.g. ``data_format='channels_first'``) raise."""
        import inspect
        cfg = dict(config)
        bis = cfg.pop("batch_input_shape", None)
        if bis is not None and "input_shape" not in cfg:
            cfg["input_shape"] = tuple(bis[1:])
        if cfg.get("data_format", "channels_last") != "channels_last":
            raise ValueError("only data_format='channels_last' is supported")
        if tuple(np.atleast_1d(cfg.get("dilation_rate", 1))) not in ((1,), (1, 1)):
            raise ValueError("dilated convolutions are not supported")
        for k in ("kernel_size", "strides", "pool_size", "input_shape"):
            if cfg.get(k) is not None:
                cfg[k] = tuple(cfg[k])
        for k in ("kernel_initializer", "bias_initializer"):
            if isinstance(cfg.get(k), dict):          # Keras spells initializers as {"class_name": ..., "config": ...}
                cfg[k] = _INITIALIZER_NAMES.get(cfg[k].get("class_name"), cfg[k].get("class_name"))
        accepted = set()
        for klass in cls.__mro__:
            if klass is object:
                continue
            accepted |= set(inspect.signature(klass.__init__).parameters)
        return cls(**{k: v for k, v in cfg.items() if k in accepted and k not in ("self", "kw", "kwargs")})

    def __repr__(self):
        return f"<{type(self).__name__} {self.name} {self.input_shape}->{self.output_shape}>"


_INITIALIZER_NAMES = {"GlorotUniform": "glorot_uniform", "Zeros": "zeros", "Ones": "ones",
                      "VarianceScaling": "glorot_uniform", "HeNormal": "he_normal", "HeUniform": "he_uniform"}
_INITIALIZER_CLASSES = {"glorot_uniform": "GlorotUniform", "zeros": "Zeros", "ones": "Ones",
                        "he_normal": "HeNormal", "he_uniform": "HeUniform"}


def _init_config(name):
    name = name if isinstance(name, str) else getattr(name, "__name__", str(name))
    return {"class_name": _INITIALIZER_CLASSES.get(name, name), "config": {}}


def layer_from_config(spec: dict):
    """``{"class_name": ..., "config": {...}}`` -> layer (recursively for Residual blocks)."""
    import sys
    cls = getattr(sys.modules[__name__], spec["class_name"], None)
    if cls is None or not (isinstance(cls, type) and issubclass(cls, Layer)):
        raise ValueError(f"unknown layer class {spec['class_name']!r}")
    return cls.from_config(spec["config"])


def layer_to_config(layer) -> dict:
    return {"class_name": type(layer).__name__, "config": layer.get_config()}


class InputLayer(Layer):
    base_name = "input"

    def __init__(self, input_shape=None, shape=None, **kw):
        super().__init__(input_shape=input_shape if input_shape is not None else shape, **kw)


class Conv2D(Layer):
    """2-D convolution, NHWC, kernel HWIO (README.md:293: ``Conv2D(32, 3, activation='relu')``)."""
    base_name = "conv2d"

    def __init__(self, filters, kernel_size, strides=(1, 1), padding="valid", activation=None,
                 use_bias=True, kernel_initializer="glorot_uniform", bias_initializer="zeros", **kw):
        super().__init__(**kw)
        self.filters = int(filters)
        self.kernel_size = _pair(kernel_size)
        self.strides = _pair(strides)
        padding = padding.lower()
        if padding not in ("valid", "same"):
            raise ValueError("padding must be 'valid' or 'same'")
        self.padding = padding
        self.activation = _norm_activation(activation)
        self.use_bias = bool(use_bias)
        self.kernel_initializer = kernel_initializer
        self.bias_initializer = bias_initializer

    def get_config(self):
        return {**self._base_config(), "filters": self.filters, "kernel_size": list(self.kernel_size),
                "strides": list(self.strides), "padding": self.padding, "data_format": "channels_last",
                "dilation_rate": [1, 1], "activation": self.activation or "linear", "use_bias": self.use_bias,
                "kernel_initializer": _init_config(self.kernel_initializer),
                "bias_initializer": _init_config(self.bias_initializer)}

    def pad_amounts(self, input_shape=None) -> tuple[int, int]:
        """Symmetric zero padding (top/left); 'same' with odd kernels and these strides."""
        if self.padding == "valid":
            return 0, 0
        H, W, _ = input_shape or self.input_shape
        out = []
        for size, k, s in ((H, self.kernel_size[0], self.strides[0]), (W, self.kernel_size[1], self.strides[1])):
            o = math.ceil(size / s)
            total = max((o - 1) * s + k - size, 0)
            out.append(total // 2)       # TF puts the extra pixel at the end; begin = total // 2
        return out[0], out[1]

    def compute_output_shape(self, input_shape):
        if len(input_shape) != 3:
            raise ValueError(f"Conv2D expects (H, W, C) inputs, got {input_shape}")
        H, W, _ = input_shape
        (kh, kw), (sh, sw) = self.kernel_size, self.strides
        if self.padding == "valid":
            oh, ow = (H - kh) // sh + 1, (W - kw) // sw + 1
        else:
            oh, ow = math.ceil(H / sh), math.ceil(W / sw)
        if oh <= 0 or ow <= 0:
            raise ValueError(f"Conv2D {self.name}: kernel {self.kernel_size} larger than input {input_shape}")
        return (oh, ow, self.filters)

    def param_specs(self):
        ci = self.input_shape[2]
        kh, kw = self.kernel_size
        specs = [ParamSpec(f"{self.name}/kernel:0", (kh, kw, ci, self.filters), self.kernel_initializer,
                           fan_in=kh * kw * ci, fan_out=kh * kw * self.filters)]
        if self.use_bias:
            specs.append(ParamSpec(f"{self.name}/bias:0", (self.filters,), self.bias_initializer))
        return specs


class MaxPooling2D(Layer):
    """Max pooling; the default (2, 2) window / stride 2 / 'valid' is what README.md:294 uses
    (fused fast path).  Other square windows / strides / 'same' padding use the general kernel."""
    base_name = "max_pooling2d"

    def __init__(self, pool_size=(2, 2), strides=None, padding="valid", **kw):
        super().__init__(**kw)
        self.pool_size = _pair(pool_size)
        self.strides = _pair(strides) if strides is not None else self.pool_size
        self.padding = padding.lower()
        if self.padding not in ("valid", "same"):
            raise ValueError("padding must be 'valid' or 'same'")
        if self.pool_size[0] != self.pool_size[1] or self.strides[0] != self.strides[1]:
            raise ValueError("MaxPooling2D supports square windows and strides")

    def get_config(self):
        return {**self._base_config(), "pool_size": list(self.pool_size), "strides": list(self.strides),
                "padding": self.padding, "data_format": "channels_last"}

    @property
    def is_default_2x2(self) -> bool:
        return self.pool_size == (2, 2) and self.strides == (2, 2) and self.padding == "valid"

    def pad_amount(self) -> int:
        if self.padding == "valid":
            return 0
        H = self.input_shape[0]
        k, s = self.pool_size[0], self.strides[0]
        total = max((math.ceil(H / s) - 1) * s + k - H, 0)
        return total // 2

    def compute_output_shape(self, input_shape):
        H, W, C = input_shape
        k, s = self.pool_size[0], self.strides[0]
        if self.padding == "valid":
            return ((H - k) // s + 1, (W - k) // s + 1, C)
        return (math.ceil(H / s), math.ceil(W / s), C)


class Flatten(Layer):
    """Row-major flatten of (H, W, C) — a view; no data moves (README.md:295)."""
    base_name = "flatten"

    def get_config(self):
        return {**self._base_config(), "data_format": "channels_last"}

    def compute_output_shape(self, input_shape):
        return (int(np.prod(input_shape)),)


class Dense(Layer):
    """Fully connected layer, kernel (in, out) (README.md:296-297)."""
    base_name = "dense"

    def __init__(self, units, activation=None, use_bias=True, kernel_initializer="glorot_uniform",
                 bias_initializer="zeros", **kw):
        super().__init__(**kw)
        self.units = int(units)
        self.activation = _norm_activation(activation)
        self.use_bias = bool(use_bias)
        self.kernel_initializer = kernel_initializer
        self.bias_initializer = bias_initializer

    def get_config(self):
        return {**self._base_config(), "units": self.units, "activation": self.activation or "linear",
                "use_bias": self.use_bias, "kernel_initializer": _init_config(self.kernel_initializer),
                "bias_initializer": _init_config(self.bias_initializer)}

    def compute_output_shape(self, input_shape):
        if len(input_shape) != 1:
            raise ValueError(f"Dense expects flat inputs, got {input_shape}; add a Flatten layer")
        return (self.units,)

    def param_specs(self):
        k = self.input_shape[0]
        specs = [ParamSpec(f"{self.name}/kernel:0", (k, self.units), self.kernel_initializer,
                           fan_in=k, fan_out=self.units)]
        if self.use_bias:
            specs.append(ParamSpec(f"{self.name}/bias:0", (self.units,), self.bias_initializer))
        return specs


class BatchNormalization(Layer):
    """Batch normalisation over the channel axis, training statistics (ResNet-50, BASELINE.json
    config 4).  Keras weight order: gamma, beta, moving_mean, moving_variance.  A following ReLU
    is folded into the same kernels."""
    base_name = "batch_normalization"

    def __init__(self, axis=-1, momentum=0.99, epsilon=1e-3, **kw):
        super().__init__(**kw)
        if axis not in (-1, 3):
            raise ValueError("BatchNormalization supports the channel (last) axis only")
        self.momentum, self.epsilon = float(momentum), float(epsilon)

    def get_config(self):
        cfg = {**self._base_config(), "axis": [3], "momentum": self.momentum, "epsilon": self.epsilon,
               "center": True, "scale": True}
        if self.activation is not None:
            cfg["activation"] = self.activation          # a folded ReLU (not a Keras key; see from_config)
        return cfg

    @classmethod
    def from_config(cls, config):
        cfg = dict(config)
        act = cfg.pop("activation", None)
        axis = cfg.get("axis", -1)
        cfg["axis"] = axis[0] if isinstance(axis, (list, tuple)) else axis
        if not (cfg.pop("center", True) and cfg.pop("scale", True)):
            raise ValueError("BatchNormalization without gamma / beta is not supported")
        layer = super().from_config(cfg)
        layer.activation = _norm_activation(act)
        return layer

    def param_specs(self):
        c = self.input_shape[-1]
        # Keras weight order.  The moving statistics are non-trainable arena slots: they are mirrored
        # variables like the rest, exported / loaded with the weights, and kept identical on every
        # replica by the same fused all-reduce that updates the trainable ones (engine/lowering.py).
        return [ParamSpec(f"{self.name}/gamma:0", (c,), "ones"), ParamSpec(f"{self.name}/beta:0", (c,), "zeros"),
                ParamSpec(f"{self.name}/moving_mean:0", (c,), "zeros", trainable=False),
                ParamSpec(f"{self.name}/moving_variance:0", (c,), "ones", trainable=False)]


class GlobalAveragePooling2D(Layer):
    base_name = "global_average_pooling2d"

    def compute_output_shape(self, input_shape):
        return (input_shape[-1],)


class Residual(Layer):
    """``relu(main(x) + shortcut(x))`` — the building block of ResNets, as one composite layer so
    that residual networks fit the Sequential container (``shortcut=None`` is the identity)."""
    base_name = "residual"

    def __init__(self, main, shortcut=None, activation="relu", **kw):
        super().__init__(**kw)
        self.main = fold_activations(main)
        self.shortcut = fold_activations(shortcut) if shortcut else []
        self.activation = _norm_activation(activation)

    def sublayers(self):
        return self.main + self.shortcut

    def get_config(self):
        return {**self._base_config(), "activation": self.activation or "linear",
                "main": [layer_to_config(l) for l in self.main],
                "shortcut": [layer_to_config(l) for l in self.shortcut]}

    @classmethod
    def from_config(cls, config):
        cfg = dict(config)
        main = [layer_from_config(c) for c in cfg.pop("main")]
        shortcut = [layer_from_config(c) for c in cfg.pop("shortcut", [])]
        bis = cfg.pop("batch_input_shape", None)
        kw = {"input_shape": tuple(bis[1:])} if bis is not None else {}
        return cls(main, shortcut or None, activation=cfg.get("activation", "relu"), name=cfg.get("name"), **kw)

    def build(self, input_shape):
        self.input_shape = tuple(int(d) for d in input_shape)
        shape = self.input_shape
        for k, l in enumerate(self.main):
            l.name = l._given_name or f"{self.name}_main{k}_{l.base_name}"
            shape = l.build(shape)
        sc = self.input_shape
        for k, l in enumerate(self.shortcut):
            l.name = l._given_name or f"{self.name}_sc{k}_{l.base_name}"
            sc = l.build(sc)
        if tuple(sc) != tuple(shape):
            raise ValueError(f"Residual {self.name}: main path {shape} and shortcut {sc} shapes differ")
        self.output_shape = tuple(shape)
        self.built = True
        return self.output_shape

    def param_specs(self):
        return [s for l in self.sublayers() for s in l.param_specs()]


def fold_activations(layers):
    """An Activation that follows a Conv2D / Dense / BatchNormalization without one becomes that layer's
    fused activation (the same rule ``Sequential.add`` applies); any other Activation stays a layer."""
    out = []
    for l in layers:
        if (type(l).__name__ in ("Activation", "ReLU") and out and getattr(out[-1], "activation", "x") is None
                and type(out[-1]).__name__ in ("Conv2D", "Dense", "BatchNormalization")):
            out[-1].activation = l.activation
        else:
            out.append(l)
    return out


def flatten_layers(layers):
    """Leaf layers in parameter order (Residual expands to main path then shortcut)."""
    out = []
    for l in layers:
        if isinstance(l, Residual):
            out.extend(flatten_layers(l.sublayers()))
        else:
            out.append(l)
    return out


class Activation(Layer):
    base_name = "activation"

    def __init__(self, activation, **kw):
        super().__init__(**kw)
        self.activation = _norm_activation(activation)

    def get_config(self):
        return {**self._base_config(), "activation": self.activation or "linear"}


class ReLU(Activation):
    base_name = "re_lu"

    def __init__(self, **kw):
        super().__init__("relu", **kw)

    def get_config(self):
        return self._base_config()

    @classmethod
    def from_config(cls, config):
        bis = config.get("batch_input_shape")
        return cls(name=config.get("name"), **({"input_shape": tuple(bis[1:])} if bis else {}))


def initialize(spec: ParamSpec, rng: np.random.Generator) -> np.ndarray:
    """Keras initialisers: glorot_uniform = U(-l, l), l = sqrt(6 / (fan_in + fan_out))."""
    if spec.init in ("zeros", "Zeros"):
        return np.zeros(spec.shape, np.float32)
    if spec.init in ("ones", "Ones"):
        return np.ones(spec.shape, np.float32)
    if spec.init in ("glorot_uniform", "GlorotUniform"):
        limit = math.sqrt(6.0 / max(1, spec.fan_in + spec.fan_out))
        return rng.uniform(-limit, limit, size=spec.shape).astype(np.float32)
    if spec.init in ("he_normal", "HeNormal"):
        std = math.sqrt(2.0 / max(1, spec.fan_in))
        return (rng.standard_normal(size=spec.shape) * std).astype(np.float32)
    raise ValueError(f"unknown initializer {spec.init!r}")
