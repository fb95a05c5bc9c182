"""Keras-style front end: shape inference, parameter layout, fit/History contract."""
import numpy as np
import pytest

import distributed_b200 as tf
from distributed_b200.engine.arena import ArenaLayout


def _readme_model():
    model = tf.keras.Sequential([
        tf.keras.layers.Conv2D(32, 3, activation='relu', input_shape=(28, 28, 1)),
        tf.keras.layers.MaxPooling2D(),
        tf.keras.layers.Flatten(),
        tf.keras.layers.Dense(64, activation='relu'),
        tf.keras.layers.Dense(10)
    ])
    model.compile(loss=tf.keras.losses.SparseCategoricalCrossentropy(from_logits=True),
                  optimizer=tf.keras.optimizers.SGD(learning_rate=0.001), metrics=['accuracy'])
    return model


def test_shapes_params_and_keras_weight_order():
    m = _readme_model()
    shapes = [l.output_shape for l in m.layers]
    assert shapes == [(26, 26, 32), (13, 13, 32), (5408,), (64,), (10,)]
    assert m.count_params() == 347146
    names = m.weight_names()
    assert names == ["conv2d/kernel:0", "conv2d/bias:0", "dense/kernel:0", "dense/bias:0",
                     "dense_1/kernel:0", "dense_1/bias:0"]                          # 6 tensors (README.md:403)
    w = m.get_weights()
    assert [a.shape for a in w] == [(3, 3, 1, 32), (32,), (5408, 64), (64,), (64, 10), (10,)]
    lim = np.sqrt(6 / (9 + 9 * 32))
    assert np.abs(w[0]).max() <= lim and np.abs(w[0]).max() > 0.5 * lim and not w[1].any()
    lay = ArenaLayout(m.layers)
    assert lay.count == 347146 and lay.n_total == lay.n_params + 4


def test_same_padding_and_strides_shape_inference():
    m = tf.keras.Sequential([tf.keras.layers.Conv2D(8, 7, strides=2, padding="same", input_shape=(224, 224, 3)),
                             tf.keras.layers.Conv2D(16, 3, padding="same")])
    assert m.layers[0].output_shape == (112, 112, 8) and m.layers[1].output_shape == (112, 112, 16)
    assert m.layers[0].pad_amounts() == (2, 2) and m.layers[1].pad_amounts() == (1, 1)


def test_fit_contract_local_cpu():
    m = _readme_model()
    rng = np.random.default_rng(0)
    x = rng.random((1000, 28, 28, 1))                      # float64, like x_train / 255.0
    y = rng.integers(0, 10, 1000)
    hist = m.fit(x, y, batch_size=64, epochs=3, steps_per_epoch=5, verbose=0)
    assert list(hist.history) == ["loss", "accuracy"] and len(hist.history["loss"]) == 3
    assert abs(hist.history["loss"][0] - np.log(10)) < 0.15          # starts near ln 10 (README.md:309)
    for a in hist.history["accuracy"]:                               # accuracy = integer / 320
        assert abs(a * 320 - round(a * 320)) < 1e-3
    assert hist.metrics is hist.history and max(hist.metrics["accuracy"]) <= 1.0


def test_training_reduces_loss_on_learnable_data():
    (x, y), _ = tf.keras.datasets.mnist.synthetic(2048, 16, seed=3)
    x = x.reshape(len(x), 28, 28, 1) / 255.0
    m = tf.keras.Sequential([tf.keras.layers.Conv2D(8, 3, activation="relu", input_shape=(28, 28, 1)),
                             tf.keras.layers.MaxPooling2D(), tf.keras.layers.Flatten(),
                             tf.keras.layers.Dense(32, activation="relu"), tf.keras.layers.Dense(10)])
    m.compile(loss=tf.keras.losses.SparseCategoricalCrossentropy(from_logits=True),
              optimizer=tf.keras.optimizers.SGD(learning_rate=0.05), metrics=["accuracy"])
    h = m.fit(x, y, batch_size=64, epochs=4, verbose=0)
    assert h.history["loss"][-1] < 0.5 * h.history["loss"][0]
    assert h.history["accuracy"][-1] > 0.8
    loss, acc = m.evaluate(x[:256], y[:256], batch_size=64)
    assert acc > 0.8


def test_fit_matches_pytorch_autograd_reference():
    """One SGD step of the engine == one SGD step of the same model written in PyTorch."""
    import torch
    import torch.nn.functional as F
    m = _readme_model()
    rng = np.random.default_rng(1)
    x = rng.random((64, 28, 28, 1)).astype(np.float32)
    y = rng.integers(0, 10, 64)
    w0 = [torch.tensor(a, requires_grad=True) for a in m.get_weights()]
    m.fit(x, y, batch_size=64, epochs=1, steps_per_epoch=1, verbose=0)
    w1 = m.get_weights()
    xt = torch.from_numpy(x).permute(0, 3, 1, 2)
    h = F.relu(F.conv2d(xt, w0[0].permute(3, 2, 0, 1), w0[1]))
    h = F.max_pool2d(h, 2).permute(0, 2, 3, 1).reshape(64, -1)          # Flatten in H, W, C order
    h = F.relu(h @ w0[2] + w0[3])
    loss = F.cross_entropy(h @ w0[4] + w0[5], torch.from_numpy(y))
    loss.backward()
    for a0, a1 in zip(w0, w1):
        ref = (a0 - 0.001 * a0.grad).detach().numpy()
        assert np.allclose(a1, ref, atol=1e-6), np.abs(a1 - ref).max()
    assert abs(m.history.history["loss"][0] - loss.item()) < 1e-4


def test_errors():
    with pytest.raises(NotImplementedError):
        tf.keras.losses.SparseCategoricalCrossentropy(from_logits=False)
    with pytest.raises(NotImplementedError):
        tf.keras.optimizers.SGD(momentum=0.9)
    m = _readme_model()
    with pytest.raises(ValueError):
        m.fit(np.zeros((10, 28, 28, 1)), np.zeros(10), batch_size=64, verbose=0)


def test_fit_refuses_what_it_would_otherwise_ignore():
    m = _readme_model()
    x, y = np.zeros((64, 28, 28, 1), np.float32), np.zeros(64)
    with pytest.raises(TypeError, match="validation_split"):
        m.fit(x, y, batch_size=64, verbose=0, validation_split=0.1)
    with pytest.raises(NotImplementedError):
        m.fit(x, y, batch_size=64, verbose=0, shuffle=True)


def test_validation_data_and_progress_lines(capsys):
    """The console contract of README.md:307-311: 'Train on N samples', 'Epoch k/n', then
    'seen/N [bar] - ETA: .. - loss: .. - accuracy: ..' (an epoch cut short by steps_per_epoch still has samples
    left, hence the ETA), metrics being streaming means over the samples seen this epoch."""
    m = _readme_model()
    rng = np.random.default_rng(0)
    x, y = rng.random((640, 28, 28, 1)).astype(np.float32), rng.integers(0, 10, 640)
    h = m.fit(x, y, batch_size=64, epochs=2, steps_per_epoch=5, validation_data=(x[:128], y[:128]))
    out = capsys.readouterr().out.splitlines()
    assert out[0] == "Train on 640 samples" and out[1] == "Epoch 1/2"
    import re
    assert re.fullmatch(r"320/640 \[=+>\.+\] - ETA: \d+s - loss: \d\.\d{4} - accuracy: 0\.\d{4}"
                        r" - val_loss: \d\.\d{4} - val_accuracy: 0\.\d{4}", out[2]), out[2]
    assert set(h.history) == {"loss", "accuracy", "val_loss", "val_accuracy"} and len(h.history["val_loss"]) == 2
    h = m.fit(x, y, batch_size=64, epochs=1)                     # whole array: the line ends with the timing instead
    last = capsys.readouterr().out.splitlines()[-1]
    assert re.fullmatch(r"640/640 \[=+\] - \d+s \d+ms/step - loss: \d\.\d{4} - accuracy: 0\.\d{4}", last), last


def test_version_and_datasets():
    assert tf.utils.tf_version() == "2.0" and tf.__version__.startswith("2.0")
    (xtr, ytr), (xte, yte) = tf.keras.datasets.mnist.load_data()
    assert xtr.shape == (60000, 28, 28) and ytr.shape == (60000,) and xte.shape == (10000, 28, 28)
    assert xtr.dtype == np.uint8 and int(ytr.max()) == 9


def test_changing_the_learning_rate_between_fits_takes_effect():
    """lr is baked into the captured step; a new value must rebuild the trainer around the same weights."""
    import numpy as np
    import distributed_b200 as tf
    rng = np.random.default_rng(0)
    x = rng.random((64, 28, 28, 1)).astype(np.float32)
    y = rng.integers(0, 10, 64)
    m = tf.models.mnist_cnn(learning_rate=0.01)
    m.fit(x, y, batch_size=64, epochs=1, steps_per_epoch=1, verbose=0)
    w1 = m.get_weights()
    m.optimizer.learning_rate = 0.0
    m.fit(x, y, batch_size=64, epochs=1, steps_per_epoch=1, verbose=0)
    assert all(np.array_equal(a, b) for a, b in zip(w1, m.get_weights()))       # lr = 0: nothing moves
    m.optimizer.learning_rate = 0.05
    m.fit(x, y, batch_size=64, epochs=1, steps_per_epoch=1, verbose=0)
    assert any(not np.array_equal(a, b) for a, b in zip(w1, m.get_weights()))


def test_learning_rate_scheduler_and_model_checkpoint(tmp_path):
    import numpy as np
    import distributed_b200 as tf
    rng = np.random.default_rng(1)
    x = rng.random((128, 28, 28, 1)).astype(np.float32)
    y = rng.integers(0, 10, 128)
    seen = []

    def schedule(epoch, lr):
        seen.append((epoch, lr))
        return 0.02 if epoch == 0 else 0.0            # second epoch must not move the weights

    m = tf.models.mnist_cnn(learning_rate=0.5)
    ck = tf.keras.callbacks.ModelCheckpoint(str(tmp_path / "w-{epoch}.npz"))
    m.fit(x, y, batch_size=64, epochs=2, steps_per_epoch=1, verbose=0,
          callbacks=[tf.keras.callbacks.LearningRateScheduler(schedule), ck])
    assert seen == [(0, 0.5), (1, 0.02)]
    a, b = np.load(tmp_path / "w-1.npz"), np.load(tmp_path / "w-2.npz")
    keys = [k for k in a.files if k.endswith("--0")]          # "conv2d/kernel:0" is stored as "conv2d__kernel--0"
    assert keys and all(np.array_equal(a[k], b[k]) for k in keys)


def test_early_stopping_stops_fit():
    import numpy as np
    import distributed_b200 as tf
    rng = np.random.default_rng(2)
    x = rng.random((64, 28, 28, 1)).astype(np.float32)
    y = rng.integers(0, 10, 64)
    m = tf.models.mnist_cnn(learning_rate=0.0)                 # the loss cannot improve
    es = tf.keras.callbacks.EarlyStopping(monitor="loss", patience=1)
    h = m.fit(x, y, batch_size=64, epochs=10, steps_per_epoch=1, verbose=0, callbacks=[es])
    assert len(h.history["loss"]) == 3 and es.stopped_epoch == 2
