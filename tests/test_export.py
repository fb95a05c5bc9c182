"""Model export (README.md:234-247): a real Keras-2 HDF5 file written without h5py, layer configs that
round-trip (incl. residual networks), and the base64 hand-off to the driver."""
import json
import struct

import numpy as np
import pytest

import distributed_b200 as tf
from distributed_b200 import export
from distributed_b200.keras import layers as L
from distributed_b200.utils import hdf5


def _compiled(model, lr=0.01):
    model.compile(loss=tf.keras.losses.SparseCategoricalCrossentropy(from_logits=True),
                  optimizer=tf.keras.optimizers.SGD(lr), metrics=["accuracy"])
    return model


def test_saved_model_is_an_hdf5_file_in_the_keras_2_layout(tmp_path):
    model = tf.models.mnist_cnn()
    path = tmp_path / "trained-0.hdf5"
    model.save(path)
    raw = path.read_bytes()
    assert raw[:8] == b"\x89HDF\r\n\x1a\n"                              # HDF5 signature, superblock version 0
    assert raw[8] == 0 and raw[13] == 8 and raw[14] == 8
    eof = struct.unpack_from("<Q", raw, 40)[0]
    assert eof == len(raw)                                              # end-of-file address == file size
    root = hdf5.read_file(path)
    cfg = json.loads(hdf5.as_text(root.attrs["model_config"]))
    assert cfg["class_name"] == "Sequential"
    assert [l["class_name"] for l in cfg["config"]["layers"]] == ["Conv2D", "MaxPooling2D", "Flatten", "Dense", "Dense"]
    assert cfg["config"]["layers"][0]["config"]["batch_input_shape"] == [None, 28, 28, 1]
    assert hdf5.as_text(root.attrs["keras_version"]).startswith("2.")
    mw = root["model_weights"]
    assert [hdf5.as_text(n) for n in mw.attrs["layer_names"]] == ["conv2d", "max_pooling2d", "flatten", "dense", "dense_1"]
    assert [hdf5.as_text(n) for n in mw["dense"].attrs["weight_names"]] == ["dense/kernel:0", "dense/bias:0"]
    assert mw["conv2d/conv2d/kernel:0"].shape == (3, 3, 1, 32) and mw["conv2d/conv2d/kernel:0"].dtype == np.float32
    assert mw["dense/dense/kernel:0"].shape == (5408, 64) and mw["dense_1/dense_1/bias:0"].shape == (10,)
    assert len(np.atleast_1d(mw["flatten"].attrs["weight_names"])) == 0
    for a, b in zip(model.get_weights(), tf.keras.models.load_model(path).get_weights()):
        assert np.array_equal(a, b)


def test_hdf5_reader_understands_a_file_written_by_libhdf5():
    """The classic on-disk structures (superblock v0, symbol-table groups, version-1 object headers and
    attributes) are shared by the writer and the reader; the reader is checked against a file produced by the
    real library (a MATLAB v7.3 file shipped with SciPy's tests: HDF5 behind a 512-byte user block)."""
    import scipy.io
    from pathlib import Path
    f = Path(scipy.io.__file__).parent / "matlab" / "tests" / "data" / "testhdf5_7.4_GLNX86.mat"
    if not f.exists():
        pytest.skip("SciPy test data not installed")
    root = hdf5.read_file(f)
    x = root["testdouble"]
    assert x.shape == (9, 1) and np.allclose(x.ravel(), np.arange(9) * np.pi / 4)


def test_weights_only_file_and_base64_roundtrip(tmp_path):
    model = tf.models.mnist_cnn()
    model.save_weights(tmp_path / "w.h5")
    root = hdf5.read_file(tmp_path / "w.h5")
    assert "model_weights" not in root and "conv2d" in root          # Keras' weights-only layout: groups at the root
    text = export.to_base64(tmp_path / "w.h5")                        # README.md:240
    export.from_base64(text, tmp_path / "model.hdf5")                 # README.md:246
    other = tf.models.mnist_cnn()
    other.set_weights([np.zeros_like(w) for w in other.get_weights()])
    other.load_weights(tmp_path / "model.hdf5")
    for a, b in zip(model.get_weights(), other.get_weights()):
        assert np.array_equal(a, b)


@pytest.mark.parametrize("ext", ["hdf5", "npz"])
def test_residual_network_save_load_predict(tmp_path, ext):
    """save() -> load_model() -> predict() on a network with Residual blocks, BatchNormalization (non-default
    momentum / epsilon, folded ReLUs), strided 'same' convolutions and general pooling."""
    layers = [L.Conv2D(8, 3, strides=2, padding="same", use_bias=False, input_shape=(16, 16, 3)),
              L.BatchNormalization(momentum=0.9, epsilon=2e-5), L.ReLU(),
              L.MaxPooling2D(3, strides=2, padding="same"),
              tf.models.bottleneck(8, 1, project=True), tf.models.bottleneck(8, 1, project=False),
              L.GlobalAveragePooling2D(), L.Dense(10)]
    model = _compiled(tf.keras.Sequential(layers))
    rng = np.random.default_rng(0)
    x = rng.standard_normal((8, 16, 16, 3)).astype(np.float32)
    y = rng.integers(0, 10, 8)
    model.fit(x, y, batch_size=8, epochs=1, steps_per_epoch=2, verbose=0)      # moves weights and moving statistics
    path = tmp_path / f"resnet.{ext}"
    model.save(path)
    loaded = tf.keras.models.load_model(path)
    assert loaded.weight_names() == model.weight_names()
    bn = [l for l in L.flatten_layers(loaded.layers) if isinstance(l, L.BatchNormalization)]
    assert bn[0].momentum == 0.9 and bn[0].epsilon == 2e-5 and bn[0].activation == "relu"
    assert json.dumps(loaded.get_config(), sort_keys=True) == json.dumps(model.get_config(), sort_keys=True)
    _compiled(loaded)
    assert np.array_equal(model.predict(x, batch_size=8), loaded.predict(x, batch_size=8))


def test_layer_configs_accept_genuine_keras_keys_and_reject_what_cannot_be_honoured():
    cfg = {"name": "conv2d", "trainable": True, "batch_input_shape": [None, 28, 28, 1], "dtype": "float32",
           "filters": 32, "kernel_size": [3, 3], "strides": [1, 1], "padding": "valid", "data_format": "channels_last",
           "dilation_rate": [1, 1], "activation": "relu", "use_bias": True,
           "kernel_initializer": {"class_name": "GlorotUniform", "config": {"seed": None}},
           "bias_initializer": {"class_name": "Zeros", "config": {}}, "kernel_regularizer": None,
           "bias_regularizer": None, "activity_regularizer": None, "kernel_constraint": None, "bias_constraint": None}
    layer = L.Conv2D.from_config(cfg)
    assert (layer.filters, layer.kernel_size, layer.activation, layer.input_shape_arg) == (32, (3, 3), "relu", (28, 28, 1))
    assert layer.kernel_initializer == "glorot_uniform" and layer.bias_initializer == "zeros"
    with pytest.raises(ValueError):
        L.Conv2D.from_config({**cfg, "data_format": "channels_first"})
    with pytest.raises(ValueError):
        L.Conv2D.from_config({**cfg, "dilation_rate": [2, 2]})
    with pytest.raises(ValueError):
        L.layer_from_config({"class_name": "LSTM", "config": {}})


def test_hdf5_roundtrip_of_random_trees():
    """Writer and reader agree on arbitrary small trees: nested groups (more entries than the default leaf size),
    float / int data sets of any rank incl. empty ones, string and numeric attributes."""
    from hypothesis import given, settings, strategies as st

    names = st.text(alphabet="abcdefghijklmnopqrstuvwxyz0123456789_:", min_size=1, max_size=12)
    shapes = st.lists(st.integers(0, 5), min_size=0, max_size=4).map(tuple)

    @st.composite
    def trees(draw, depth=0):
        g = {"attrs": {}, "children": {}}
        for k in draw(st.lists(names, max_size=3, unique=True)):
            g["attrs"][k] = draw(st.one_of(st.text(alphabet="abc xyz{}:,\"", max_size=30),
                                           st.lists(names, min_size=1, max_size=4),
                                           st.integers(-5, 5), st.floats(-1e3, 1e3, width=32)))
        for k in draw(st.lists(names, max_size=12 if depth == 0 else 3, unique=True)):
            if depth < 2 and draw(st.booleans()):
                g["children"][k] = draw(trees(depth + 1))
            else:
                shape = draw(shapes)
                dt = draw(st.sampled_from([np.float32, np.float64, np.int32, np.int64, np.uint8]))
                g["children"][k] = (np.arange(int(np.prod(shape)), dtype=np.float64).reshape(shape) - 3).astype(dt)
        return g

    def build(spec):
        g = hdf5.Group()
        g.attrs.update(spec["attrs"])
        for k, v in spec["children"].items():
            if isinstance(v, dict):
                g.children[k] = build(v)
            else:
                g.create_dataset(k, v)
        return g

    def same(spec, g):
        assert set(spec["children"]) == set(g.children)
        for k, v in spec["attrs"].items():
            got = g.attrs[k]
            if isinstance(v, str):
                assert hdf5.as_text(got) == v
            elif isinstance(v, list):
                assert [hdf5.as_text(x) for x in np.atleast_1d(got)] == v
            else:
                assert float(got) == float(np.float32(v)) or float(got) == float(v)
        for k, v in spec["children"].items():
            if isinstance(v, dict):
                same(v, g.children[k])
            else:
                assert g.children[k].dtype == v.dtype and g.children[k].shape == v.shape
                assert np.array_equal(g.children[k], v)

    import tempfile, os

    @settings(max_examples=60, deadline=None, derandomize=True)      # the same trees on every run
    @given(trees())
    def run(spec):
        fd, path = tempfile.mkstemp(suffix=".h5")
        os.close(fd)
        try:
            hdf5.write_file(path, build(spec))
            assert hdf5.is_hdf5(path)
            same(spec, hdf5.read_file(path))
        finally:
            os.remove(path)

    run()
