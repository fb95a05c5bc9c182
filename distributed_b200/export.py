"""Model and weight export in Keras layer order (README.md:234-247: ``save_model_hdf5`` then base64).

Tensors are written in the order and layouts Keras uses — for the README model
``conv2d/kernel:0`` (3,3,1,32) HWIO, ``conv2d/bias:0``, ``dense/kernel:0`` (5408,64) (in,out)
with Flatten in H,W,C order, ``dense/bias:0``, ``dense_1/kernel:0`` (64,10), ``dense_1/bias:0``.

Formats, chosen by the file extension:

* ``.h5`` / ``.hdf5`` — a real HDF5 file in the Keras-2 layout, written by the
  dependency-free writer in ``utils/hdf5.py`` (h5py / libhdf5 are not needed): root attributes
  ``model_config`` (JSON, ``{"class_name": "Sequential", "config": ...}``), ``keras_version``, ``backend``;
  group ``model_weights`` with attribute ``layer_names``; one group per layer with attribute
  ``weight_names``; one float32 data set per weight at ``model_weights/<layer>/<weight name>``.
  ``save_weights`` writes the same groups at the root (Keras' weights-only layout).
* anything else (``.npz``) — a NumPy archive with the same names, order and config.

``load_weights`` / ``load_model`` recognise either container by its magic bytes, so the README's
save -> base64 -> decode -> load round trip works with both.
"""
from __future__ import annotations

import base64
import json
from pathlib import Path

import numpy as np

from .utils import hdf5

KERAS_VERSION = "2.2.4-tf"          # what TensorFlow 2.0 (the reference's version, README.md:260) writes


def weights_dict(model) -> dict[str, np.ndarray]:
    return dict(zip(model.weight_names(), model.get_weights()))


# ---- npz ---------------------------------------------------------------------------------------
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


# ---- HDF5 (Keras-2 layout) -----------------------------------------------------------------------
def _save_h5(path, model, tensors, config):
    root = hdf5.Group()
    g = root.create_group("model_weights") if config is not None else root
    if config is not None:
        root.attrs["model_config"] = json.dumps({"class_name": "Sequential", "config": config},
                                                separators=(",", ":"))
        root.attrs["keras_version"] = KERAS_VERSION
        root.attrs["backend"] = "tensorflow"
    g.attrs["layer_names"] = [l.name for l in model.layers]
    g.attrs["backend"] = "tensorflow"
    g.attrs["keras_version"] = KERAS_VERSION
    for l in model.layers:
        lg = g.create_group(l.name)
        names = [s.name for s in l.param_specs()]
        lg.attrs["weight_names"] = names
        for n in names:
            node = lg
            *dirs, leaf = n.split("/")                  # "conv2d/kernel:0" -> group conv2d, data set kernel:0
            for d in dirs:
                node = node.children[d] if d in node.children else node.create_group(d)
            node.create_dataset(leaf, np.asarray(tensors[n], np.float32))
    hdf5.write_file(path, root)


def _load_h5(path):
    root = hdf5.read_file(path)
    config = None
    if "model_config" in root.attrs:
        config = json.loads(hdf5.as_text(root.attrs["model_config"]))["config"]
        if isinstance(config, list):                    # very old Keras: the config IS the layer list
            config = {"name": "sequential", "layers": config}
    g = root["model_weights"] if "model_weights" in root else root
    tensors = {}
    for ln in np.atleast_1d(g.attrs["layer_names"]):
        lg = g[hdf5.as_text(ln)]
        for wn in np.atleast_1d(lg.attrs.get("weight_names", [])):
            wn = hdf5.as_text(wn)
            tensors[wn] = np.asarray(lg[wn], np.float32)
    return tensors, config


def _write(model, path, config):
    path = Path(path)
    tensors = weights_dict(model)
    if path.suffix in (".h5", ".hdf5"):
        _save_h5(path, model, tensors, config)
    else:
        _save_npz(path, tensors, config)


def save_weights(model, path):
    _write(model, path, None)


def save_model(model, path):
    _write(model, path, model.get_config())


def _read(path):
    path = Path(path)
    with open(path, "rb") as f:
        magic = f.read(8)
    if magic[:2] == b"PK":
        return _load_npz(path)
    return _load_h5(path)


def load_weights(model, path):
    tensors, _ = _read(path)
    names = model.weight_names()
    missing = [n for n in names if n not in tensors]
    if missing:
        raise ValueError(f"weights file lacks {missing}")
    model.set_weights([tensors[n] for n in names])


def load_model(path):
    from .keras.models import Sequential
    tensors, config = _read(path)
    if config is None:
        raise ValueError("file holds weights only; build the model and call load_weights")
    model = Sequential.from_config(config)
    model.set_weights([tensors[n] for n in model.weight_names()])
    return model


def to_base64(path) -> str:
    """``base64enc::base64encode(model_file)`` — how worker 0 ships its file (README.md:240)."""
    return base64.b64encode(Path(path).read_bytes()).decode()


def from_base64(text: str, path) -> None:
    """``write(base64decode(...), "model.hdf5")`` on the driver (README.md:246)."""
    Path(path).write_bytes(base64.b64decode(text))
