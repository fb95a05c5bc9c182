"""Callbacks.  The reference passes none and TensorFlow warns about the missing
``ModelCheckpoint`` (README.md:400); it is provided here so a run can actually be resumed."""
from __future__ import annotations


class Callback:
    model = None

    def set_model(self, model):
        self.model = model

    def on_train_begin(self, logs=None): pass
    def on_train_end(self, logs=None): pass
    def on_epoch_begin(self, epoch, logs=None): pass
    def on_epoch_end(self, epoch, logs=None): pass


class ModelCheckpoint(Callback):
    """Saves the weights (Keras layer order) after every epoch; only the chief writes."""

    def __init__(self, filepath, save_weights_only=True, save_freq="epoch", chief_only=True):
        self.filepath = str(filepath)
        self.chief_only = chief_only

    def on_epoch_end(self, epoch, logs=None):
        if self.chief_only and self.model._strategy.rank != 0:
            return
        path = self.filepath.format(epoch=epoch + 1, **(logs or {}))
        self.model.save_weights(path)


class LearningRateScheduler(Callback):
    """``schedule(epoch, lr) -> lr`` evaluated at the start of every epoch (Keras signature).  The learning
    rate is a constant of the captured step graph, so a changed value re-captures the step (a few
    milliseconds) around the current weights; every replica evaluates the same schedule."""

    def __init__(self, schedule, verbose=0):
        self.schedule, self.verbose = schedule, verbose

    def on_epoch_begin(self, epoch, logs=None):
        opt = self.model.optimizer
        lr = float(self.schedule(epoch, float(opt.learning_rate)))
        if self.verbose:
            print(f"Epoch {epoch + 1}: LearningRateScheduler setting learning rate to {lr}.")
        opt.learning_rate = lr


class EarlyStopping(Callback):
    """Stops ``fit`` when ``monitor`` has not improved by ``min_delta`` for ``patience`` epochs.  The metric is
    the cross-replica value, identical on every worker, so all replicas stop on the same epoch."""

    def __init__(self, monitor="loss", min_delta=0.0, patience=0, mode="auto"):
        self.monitor, self.min_delta, self.patience = monitor, abs(float(min_delta)), int(patience)
        self.maximise = mode == "max" or (mode == "auto" and "acc" in monitor)

    def on_train_begin(self, logs=None):
        self.best, self.wait, self.stopped_epoch = None, 0, None

    def on_epoch_end(self, epoch, logs=None):
        v = (logs or {}).get(self.monitor)
        if v is None:
            return
        better = self.best is None or (v > self.best + self.min_delta if self.maximise
                                       else v < self.best - self.min_delta)
        if better:
            self.best, self.wait = v, 0
            return
        self.wait += 1
        if self.wait > self.patience:
            self.stopped_epoch = epoch
            self.model.stop_training = True

