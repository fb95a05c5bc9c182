"""Weight export in Keras layer order (README.md:234-247: ``save_model_hdf5`` then base64).

Tensors are written in the order and layouts Keras uses — for the README model
``conv2d/kernel:0`` (3,3,1,32) HWIO, ``conv2d/bias:0``, ``dense/kernel:0`` (5408,64) (in,out)
with Flatten in H,W,C order, ``dense/bias:0``, ``dense_1/kernel:0`` (64,10), ``dense_1/bias:0``.

Formats: ``.npz`` is always available.  ``.h5`` / ``.hdf5`` uses h5py when it is installed and
writes the Keras 2 weight-file layout; without h5py (this image) the same tensors are written
as an ``.npz`` payload under the requested file name, which ``load_weights`` recognises by
its magic bytes — so the README's save → base64 → decode → load round trip works either way.
"""
from __future__ import annotations

import base64
import io
import json
from pathlib import Path

import numpy as np


def _have_h5py():
    try:
        import h5py  # noqa: F401
        return True
    except Exception:
        return False


def weights_dict(model) -> dict[str, np.ndarray]:
    return dict(zip(model.weight_names(), model.get_weights()))


def _save_npz(path, tensors: dict, config: dict | None):
    payload = {k.replace("/", "__").replace(":", "--"): v for k, v in tensors.items()}
    payload["__order__"] = np.array(list(tensors.keys()))
    if config is not None:
        payload["__config__"] = np.array(json.dumps(config))
    with open(path, "wb") as f:
        np.savez(f, **payload)


def _load_npz(path):
    with np.load(path, allow_pickle=False) as f:
        order = [str(k) for k in f["__order__"]]
        tensors = {k: f[k.replace("/", "__").replace(":", "--")] for k in order}
        config = json.loads(str(f["__config__"])) if "__config__" in f.files else None
    return tensors, config


def _save_h5(path, model, tensors, config):
    import h5py
    with h5py.File(path, "w") as f:
        g = f.create_group("model_weights") if config is not None else f
        if config is not None:
            f.attrs["model_config"] = json.dumps({"class_name": "Sequential", "config": config})
        g.attrs["layer_names"] = [l.name.encode() for l in model.layers]
        g.attrs["backend"] = b"distributed_b200"
        for l in model.layers:
            lg = g.create_group(l.name)
            names = [s.name for s in l.param_specs()]
            lg.attrs["weight_names"] = [n.encode() for n in names]
            for n in names:
                lg.create_dataset(n, data=tensors[n])


def _load_h5(path):
    import h5py
    tensors, config = {}, None
    with h5py.File(path, "r") as f:
        if "model_config" in f.attrs:
            config = json.loads(f.attrs["model_config"])["config"]
        g = f["model_weights"] if "model_weights" in f else f
        for ln in g.attrs["layer_names"]:
            ln = ln.decode() if isinstance(ln, bytes) else ln
            for wn in g[ln].attrs["weight_names"]:
                wn = wn.decode() if isinstance(wn, bytes) else wn
                tensors[wn] = np.array(g[ln][wn])
    return tensors, config


def _is_zip(path) -> bool:
    with open(path, "rb") as f:
        return f.read(2) == b"PK"


def _write(model, path, config):
    path = Path(path)
    tensors = weights_dict(model)
    if path.suffix in (".h5", ".hdf5") and _have_h5py():
        _save_h5(path, model, tensors, config)
    else:
        _save_npz(path, tensors, config)


def save_weights(model, path):
    _write(model, path, None)


def save_model(model, path):
    _write(model, path, model.get_config())


def _read(path):
    path = Path(path)
    if _is_zip(path):
        return _load_npz(path)
    if _have_h5py():
        return _load_h5(path)
    raise RuntimeError(f"{path} is an HDF5 file but h5py is not installed")


def load_weights(model, path):
    tensors, _ = _read(path)
    names = model.weight_names()
    missing = [n for n in names if n not in tensors]
    if missing:
        raise ValueError(f"weights file lacks {missing}")
    model.set_weights([tensors[n] for n in names])


def load_model(path):
    from .keras import layers as L
    from .keras.models import Sequential
    tensors, config = _read(path)
    if config is None:
        raise ValueError("file holds weights only; build the model and call load_weights")
    model = Sequential(name=config.get("name", "sequential"))
    for lc in config["layers"]:
        cls = getattr(L, lc["class_name"])
        cfg = dict(lc["config"])
        for k in ("kernel_size", "strides", "pool_size", "input_shape"):
            if k in cfg and cfg[k] is not None:
                cfg[k] = tuple(cfg[k])
        model.add(cls(**cfg))
    model.set_weights([tensors[n] for n in model.weight_names()])
    return model


def to_base64(path) -> str:
    """``base64enc::base64encode(model_file)`` — how worker 0 ships its file (README.md:240)."""
    return base64.b64encode(Path(path).read_bytes()).decode()


def from_base64(text: str, path) -> None:
    """``write(base64decode(...), "model.hdf5")`` on the driver (README.md:246)."""
    Path(path).write_bytes(base64.b64decode(text))
