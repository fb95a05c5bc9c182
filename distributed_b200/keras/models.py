"""``Sequential`` with ``compile`` / ``fit`` / ``History`` (README.md:292-304, 375-392).

Observable contract taken from the reference's console output:

* ``fit(x, y, batch_size, epochs, steps_per_epoch)`` takes in-memory arrays; ``batch_size``
  is the GLOBAL batch under a strategy (README.md:367, 413: 1280 = 5 x 256) and each replica
  consumes a disjoint ``batch_size / num_workers`` slice of every global batch.
* "Train on N samples", "Epoch k/n", then ``seen/N ... - loss: - accuracy:`` where both
  metrics are streaming means over the samples seen this epoch, aggregated over all replicas
  (README.md:307-311, 413-415); every worker reports the identical value (README.md:229-231).
* ``fit`` returns a ``History`` whose ``history['accuracy']`` is a per-epoch list
  (README.md:218-220 reads ``result$metrics$accuracy``).
"""
from __future__ import annotations

import sys
import time

import numpy as np
import torch

from . import layers as L
from . import losses as losses_mod
from . import metrics as metrics_mod
from . import optimizers as optimizers_mod


class History:
    def __init__(self):
        self.history: dict[str, list[float]] = {}
        self.epoch: list[int] = []
        self.params: dict = {}

    @property
    def metrics(self):          # R: result$metrics$accuracy (README.md:220)
        return self.history

    def _append(self, epoch: int, logs: dict):
        self.epoch.append(epoch)
        for k, v in logs.items():
            self.history.setdefault(k, []).append(float(v))


class Model:
    pass


class Sequential(Model):
    def __init__(self, layers=None, name: str = "sequential"):
        from ..parallel.strategy import get_strategy
        self.name = name
        self.layers: list[L.Layer] = []
        self._strategy = get_strategy()     # variables are created under the enclosing scope()
        self.optimizer = None
        self.loss = None
        self.metrics_names: list[str] = []
        self._trainer = None
        self._pending_weights = None
        self._global_step = 0
        self.stop_training = False
        for layer in layers or []:
            self.add(layer)

    # ---- construction ----------------------------------------------------------------
    def add(self, layer: L.Layer):
        if not isinstance(layer, L.Layer):
            raise TypeError("Sequential.add expects a Layer")
        if isinstance(layer, L.Activation) and self.layers and self.layers[-1].activation is None \
                and isinstance(self.layers[-1], (L.Conv2D, L.Dense, L.BatchNormalization)):
            self.layers[-1].activation = layer.activation     # fold into the producer's epilogue
            self._rebuild_shapes()
            return
        counts: dict[str, int] = {}
        for l in self.layers:
            counts[l.base_name] = counts.get(l.base_name, 0) + 1
        if layer._given_name is None:
            k = counts.get(layer.base_name, 0)
            layer.name = layer.base_name if k == 0 else f"{layer.base_name}_{k}"
        self.layers.append(layer)
        self._trainer = None
        self._rebuild_shapes()

    def _rebuild_shapes(self):
        shape = None
        for layer in self.layers:
            if layer.input_shape_arg is not None:
                shape = layer.input_shape_arg
            if shape is None:
                return
            shape = layer.build(shape)

    @property
    def built(self) -> bool:
        return bool(self.layers) and all(l.built for l in self.layers)

    @property
    def input_shape(self):
        return (None, *self.layers[0].input_shape) if self.built else None

    @property
    def output_shape(self):
        return (None, *self.layers[-1].output_shape) if self.built else None

    def build(self, input_shape):
        shape = tuple(input_shape)[1:] if len(input_shape) == 4 or input_shape[0] is None else tuple(input_shape)
        self.layers[0].input_shape_arg = shape
        self._rebuild_shapes()

    def count_params(self) -> int:
        return sum(l.count_params() for l in self.layers)

    @property
    def weights(self):
        return [s for l in self.layers for s in l.param_specs()]

    @property
    def trainable_weights(self):
        return [s for s in self.weights if s.trainable]

    @property
    def non_trainable_weights(self):
        return [s for s in self.weights if not s.trainable]

    def summary(self, print_fn=print):
        print_fn(f'Model: "{self.name}"')
        print_fn("_" * 65)
        print_fn(f"{'Layer (type)':<29}{'Output Shape':<26}{'Param #':<10}")
        print_fn("=" * 65)
        for l in self.layers:
            shp = str((None, *l.output_shape)) if l.built else "?"
            print_fn(f"{l.name + ' (' + type(l).__name__ + ')':<29}{shp:<26}{l.count_params():<10}")
        print_fn("=" * 65)
        print_fn(f"Total params: {self.count_params():,}")
        trainable = sum(l.count_trainable_params() for l in self.layers)
        print_fn(f"Trainable params: {trainable:,}")
        print_fn(f"Non-trainable params: {self.count_params() - trainable:,}")

    # ---- compile -----------------------------------------------------------------------
    def compile(self, optimizer="sgd", loss=None, metrics=None, **kwargs):
        self.optimizer = optimizers_mod.get(optimizer)
        self.loss = losses_mod.get(loss)
        self.metrics_names = ["loss"] + metrics_mod.normalize(metrics)
        self._trainer = None

    # ---- training ------------------------------------------------------------------------
    def _ensure_trainer(self, batch_per_replica: int, sample_shape):
        if not self.built:
            self.layers[0].input_shape_arg = tuple(sample_shape)
            self._rebuild_shapes()
        if tuple(sample_shape) != tuple(self.layers[0].input_shape):
            raise ValueError(f"input samples have shape {tuple(sample_shape)}, model expects "
                             f"{self.layers[0].input_shape}")
        t = self._trainer
        # the learning rate is baked into the captured step (update kernel, BatchNorm pseudo-gradients): a
        # changed optimizer.learning_rate rebuilds the trainer around the current weights
        if t is None or t.batch != batch_per_replica or t.lr != float(self.optimizer.learning_rate):
            weights = t.get_weights() if t is not None else self._pending_weights
            if t is not None:
                t.close()
            from ..engine.trainer import Trainer
            self._trainer = Trainer(self, self._strategy, batch_per_replica, self.layers[0].input_shape)
            if weights is not None:
                self._trainer.set_weights(weights)
            self._pending_weights = None
        return self._trainer

    @staticmethod
    def _as_host_tensors(x, y, n_rows, pin):
        """fp32 NHWC samples and int32 labels for the first ``n_rows`` rows (pinned for CUDA).

        Tensors that are already fp32 / int32, contiguous and (for CUDA) pinned are used in
        place, so a caller who keeps the data set in pinned memory pays no conversion per fit.
        """
        if isinstance(x, torch.Tensor) and isinstance(y, torch.Tensor) and x.dtype == torch.float32 \
                and y.dtype == torch.int32 and x.dim() == 4 and x.is_contiguous() and y.is_contiguous() \
                and not x.is_cuda and (not pin or (x.is_pinned() and y.is_pinned())):
            return x[:n_rows], y[:n_rows]
        xt = torch.as_tensor(np.ascontiguousarray(np.asarray(x[:n_rows]), dtype=np.float32))
        if xt.dim() == 3:
            xt = xt.unsqueeze(-1)
        yt = torch.as_tensor(np.ascontiguousarray(np.asarray(y[:n_rows]).reshape(-1), dtype=np.int32))
        if pin:
            xt, yt = xt.pin_memory(), yt.pin_memory()
        return xt, yt

    def fit(self, x=None, y=None, batch_size=None, epochs=1, verbose=1, steps_per_epoch=None,
            callbacks=None, shuffle=False, initial_epoch=0, validation_data=None, **kwargs):
        """Train from in-memory arrays (README.md:304, 392).  ``batch_size`` is the GLOBAL batch under a
        strategy.  Batches are taken in order (``shuffle=True`` is refused rather than ignored);
        ``validation_data=(x, y)`` adds ``val_loss`` / ``val_accuracy`` to the logs after every epoch; any
        other keyword raises instead of being silently dropped."""
        if kwargs:
            raise TypeError(f"fit() got unsupported keyword arguments {sorted(kwargs)}; supported: x, y, batch_size, "
                            "epochs, verbose, steps_per_epoch, callbacks, shuffle=False, initial_epoch, validation_data")
        if shuffle:
            raise NotImplementedError("shuffle=True is not supported: every replica reads its slice of consecutive "
                                      "global batches; shuffle the arrays before calling fit()")
        if self.optimizer is None or self.loss is None:
            raise RuntimeError("call compile(loss=..., optimizer=...) before fit()")
        strat = self._strategy
        n_workers = strat.num_replicas_in_sync
        n_samples = len(x)
        global_batch = int(batch_size) if batch_size is not None else 32 * n_workers
        if global_batch % n_workers:
            raise ValueError(f"batch_size={global_batch} is not divisible by num_workers={n_workers}; "
                             "the reference uses batch_size = 64 * num_workers (README.md:366-367)")
        per_replica = global_batch // n_workers
        batches_in_data = n_samples // global_batch
        if batches_in_data == 0:
            raise ValueError("fewer samples than one global batch")
        steps = int(steps_per_epoch) if steps_per_epoch is not None else batches_in_data
        sample_shape = tuple(np.shape(x)[1:])
        if len(sample_shape) == 2:
            sample_shape = (*sample_shape, 1)
        trainer = self._ensure_trainer(per_replica, sample_shape)
        total_steps = steps * (epochs - initial_epoch)
        rows_needed = min(batches_in_data, self._global_step + total_steps) * global_batch
        xt, yt = self._as_host_tensors(x, y, rows_needed, pin=trainer.backend_name == "cuda")
        steps_in_src = rows_needed // global_batch
        callbacks = list(callbacks or [])
        if n_workers > 1 and not any(type(c).__name__ == "ModelCheckpoint" for c in callbacks) and verbose:
            from ..parallel.strategy import log
            log.warning("ModelCheckpoint callback is not provided. Workers will need to restart "
                        "training if any fails.")
        self.stop_training = False
        history = History()
        history.params = {"epochs": epochs, "steps": steps, "batch_size": global_batch, "samples": n_samples}
        out = sys.stdout
        # verbose=1 on a terminal: the Keras progress line is refreshed while the epoch runs (README.md:307-311
        # shows lines captured mid-epoch: "192/60000 [...] - ETA: 29s - loss: 2.3059 - accuracy: 0.2240").  The
        # epoch is then run in up to 10 chunks, each ending with a read of the running sums.
        live = verbose == 1 and hasattr(out, "isatty") and out.isatty()
        if verbose:
            print(f"Train on {n_samples} samples", file=out)
        for cb in callbacks:
            cb.set_model(self)
            cb.on_train_begin()
        for epoch in range(initial_epoch, epochs):
            if verbose:
                print(f"Epoch {epoch + 1}/{epochs}", file=out)
            for cb in callbacks:
                cb.on_epoch_begin(epoch)
            trainer = self._ensure_trainer(per_replica, sample_shape)     # a callback may have changed the lr
            trainer.reset_metrics()
            t0 = time.perf_counter()
            done = 0
            chunk = max(1, -(-steps // 10)) if live else steps
            while done < steps:
                n = min(chunk, steps - done)
                trainer.run_steps(xt, yt, self._global_step + done, steps_in_src, n, global_batch)
                done += n
                if live and done < steps:
                    loss_sum, correct, seen = trainer.metric_sums()
                    out.write("\r" + self._progress_line(seen, n_samples, time.perf_counter() - t0, done, loss_sum,
                                                          correct, final=False))
                    out.flush()
            loss_sum, correct, seen = trainer.metric_sums()
            dt = time.perf_counter() - t0
            self._global_step += steps
            logs = {"loss": loss_sum / max(seen, 1.0)}
            if "accuracy" in self.metrics_names:
                # float32 division so the value matches Keras' float32 metric (README.md:229-231)
                logs["accuracy"] = float(np.float32(correct) / np.float32(max(seen, 1.0)))
            if validation_data is not None:
                val = self.evaluate(validation_data[0], validation_data[1], batch_size=per_replica)
                val = val if isinstance(val, list) else [val]
                logs["val_loss"] = val[0]
                if len(val) > 1:
                    logs["val_accuracy"] = val[1]
            history._append(epoch, logs)
            if verbose == 2:      # Keras' one-line-per-epoch mode: no bar
                print(f"{int(seen)}/{n_samples} - {dt:.0f}s - " + " - ".join(f"{k}: {v:.4f}" for k, v in logs.items()),
                      file=out)
            elif verbose:
                line = self._progress_line(seen, n_samples, dt, steps, loss_sum, correct, final=True)
                extra = "".join(f" - {k}: {v:.4f}" for k, v in logs.items() if k.startswith("val_"))
                print(("\r" if live else "") + line + extra, file=out)
            for cb in callbacks:
                cb.on_epoch_end(epoch, logs)
            if self.stop_training:
                break
        for cb in callbacks:
            cb.on_train_end()
        self.history = history
        return history

    def _progress_line(self, seen, n_samples, dt, steps_done, loss_sum, correct, final: bool) -> str:
        """One Keras-style progress line: ``seen/total [===>....] - ETA: m:ss - loss: - accuracy:`` while samples
        remain (also at the end of an epoch cut short by ``steps_per_epoch``, exactly as README.md:309-311 and
        413-415 show), ``- 9s 23ms/step`` once the whole array has been consumed.  Both metrics are streaming
        means over the samples seen so far this epoch, summed over all replicas."""
        seen_i = int(seen)
        width = 30
        filled = min(width, int(width * seen / max(n_samples, 1)))
        bar = "=" * filled if filled >= width else "=" * max(filled - 1, 0) + (">" if filled else "") \
            + "." * (width - max(filled, 0))
        bar = (bar + "." * width)[:width]
        parts = [f"{seen_i:>{len(str(n_samples))}}/{n_samples} [{bar}]"]
        if seen_i < n_samples:
            eta = (n_samples - seen) * dt / max(seen, 1.0)
            parts.append("ETA: " + (f"{int(eta) // 60}:{int(eta) % 60:02d}" if eta >= 60 else f"{eta:.0f}s"))
        else:
            parts.append(f"{dt:.0f}s {1e3 * dt / max(steps_done, 1):.0f}ms/step")
        parts.append(f"loss: {loss_sum / max(seen, 1.0):.4f}")
        if "accuracy" in self.metrics_names:
            parts.append(f"accuracy: {float(np.float32(correct) / np.float32(max(seen, 1.0))):.4f}")
        return " - ".join(parts)

    # ---- inference (not exercised by the reference; provided for the local validation flow) --
    def predict(self, x, batch_size=None, verbose=0):
        x = np.asarray(x)
        sample_shape = tuple(x.shape[1:]) if x.ndim == 4 else (*x.shape[1:], 1)
        t = self._trainer or self._ensure_trainer(int(batch_size or 32), sample_shape)
        B = t.batch
        outs = []
        for lo in range(0, len(x), B):
            chunk = np.ascontiguousarray(x[lo:lo + B], dtype=np.float32).reshape(-1, *sample_shape)
            n = len(chunk)
            if n < B:
                chunk = np.concatenate([chunk, np.zeros((B - n, *sample_shape), np.float32)])
            outs.append(t.forward_logits(torch.from_numpy(chunk))[:n].numpy())
        return np.concatenate(outs) if outs else np.zeros((0, self.layers[-1].output_shape[-1]), np.float32)

    def evaluate(self, x, y, batch_size=None, verbose=0):
        logits = self.predict(x, batch_size)
        y = np.asarray(y).reshape(-1)[:len(logits)].astype(np.int64)
        m = logits.max(1, keepdims=True)
        lse = (m + np.log(np.exp(logits - m).sum(1, keepdims=True))).reshape(-1)
        loss = float(np.mean(lse - logits[np.arange(len(y)), y]))
        acc = float(np.mean(logits.argmax(1) == y))
        return [loss, acc] if "accuracy" in self.metrics_names else loss

    # ---- weights / export ----------------------------------------------------------------
    def get_weights(self):
        if self._trainer is not None:
            return self._trainer.get_weights()
        if self._pending_weights is not None:
            return [np.array(w) for w in self._pending_weights]
        if not self.built:
            raise RuntimeError("the model has not been built (pass input_shape to the first layer)")
        from ..engine.arena import ArenaLayout
        lay = ArenaLayout(self.layers)
        flat = lay.initial_values(self.layers, self._strategy.seed)
        return [flat[s.offset:s.offset + s.size].reshape(s.shape).copy() for s in lay.slots]

    def set_weights(self, weights):
        if self._trainer is not None:
            self._trainer.set_weights(weights)
        else:
            self._pending_weights = [np.asarray(w, np.float32) for w in weights]

    def weight_names(self) -> list[str]:
        return [s.name for l in self.layers for s in l.param_specs()]

    def save_weights(self, filepath, overwrite=True):
        from .. import export
        export.save_weights(self, filepath)

    def load_weights(self, filepath):
        from .. import export
        export.load_weights(self, filepath)

    def save(self, filepath, overwrite=True, include_optimizer=True):
        """Weights + architecture; ``save_model_hdf5`` in the README (README.md:237-238)."""
        from .. import export
        export.save_model(self, filepath)

    def get_config(self):
        """Keras-2 Sequential config: every layer as {"class_name", "config"} (Residual blocks recursively)."""
        return {"name": self.name, "layers": [L.layer_to_config(l) for l in self.layers]}

    @classmethod
    def from_config(cls, config):
        model = cls(name=config.get("name", "sequential"))
        for spec in config["layers"]:
            model.add(L.layer_from_config(spec))
        return model
