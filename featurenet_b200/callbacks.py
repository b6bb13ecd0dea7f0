"""Host-side callbacks with the Keras 2.2.4 semantics that helpers.train_model relies on.

/root/reference/helpers.py:72-97 installs, in this order, ReduceLROnPlateau(factor sqrt(0.1),
patience 5, min_lr 0.5e-6), LearningRateScheduler(lr_schedule), EarlyStopping(val_acc, max,
min_delta 0.01, patience 20) and ModelCheckpoint(val_loss, min, save_best_only).  They are
plain scalars-on-the-host logic; the only device work is ModelCheckpoint, which here is a
device-to-device snapshot of the weight arenas instead of an HDF5 file
(the reference's saved_models/temp_model.h5 collides when candidates run concurrently).
"""
import numpy as np


class Callback(object):
    def __init__(self):
        self.model = None
        self.params = {}

    def set_model(self, model):
        self.model = model

    def set_params(self, params):
        self.params = params

    def on_train_begin(self, logs=None):
        pass

    def on_train_end(self, logs=None):
        pass

    def on_epoch_begin(self, epoch, logs=None):
        pass

    def on_epoch_end(self, epoch, logs=None):
        pass

    def on_batch_begin(self, batch, logs=None):
        pass

    def on_batch_end(self, batch, logs=None):
        pass


class History(Callback):
    def on_train_begin(self, logs=None):
        self.epoch = []
        self.history = {}

    def on_epoch_end(self, epoch, logs=None):
        self.epoch.append(epoch)
        for k, v in (logs or {}).items():
            self.history.setdefault(k, []).append(v)


class LearningRateScheduler(Callback):
    def __init__(self, schedule, verbose=0):
        super(LearningRateScheduler, self).__init__()
        self.schedule = schedule

    def on_epoch_begin(self, epoch, logs=None):
        lr = float(self.model.optimizer.lr)
        try:                        # Keras 2.2.x tries (epoch, lr) first, then the old one-argument form
            lr = self.schedule(epoch, lr)
        except TypeError:
            lr = self.schedule(epoch)
        self.model.optimizer.lr = float(lr)

    def on_epoch_end(self, epoch, logs=None):
        if logs is not None:
            logs["lr"] = float(self.model.optimizer.lr)


class ReduceLROnPlateau(Callback):
    def __init__(self, monitor="val_loss", factor=0.1, patience=10, verbose=0, mode="auto", min_delta=1e-4,
                 cooldown=0, min_lr=0, **_):
        super(ReduceLROnPlateau, self).__init__()
        if factor >= 1.0:
            raise ValueError("ReduceLROnPlateau does not support a factor >= 1.0.")
        self.monitor, self.factor, self.patience = monitor, float(factor), patience
        self.min_delta, self.cooldown, self.min_lr = min_delta, cooldown, min_lr
        self.mode = "max" if (mode == "max" or (mode == "auto" and "acc" in monitor)) else "min"
        self._reset()

    def _reset(self):
        self.best = -np.inf if self.mode == "max" else np.inf
        self.cooldown_counter = 0
        self.wait = 0

    def _improved(self, current):
        if self.mode == "max":
            return current > self.best + self.min_delta
        return current < self.best - self.min_delta

    def on_train_begin(self, logs=None):
        self._reset()

    def on_epoch_end(self, epoch, logs=None):
        logs = logs if logs is not None else {}
        logs["lr"] = float(self.model.optimizer.lr)
        current = logs.get(self.monitor)
        if current is None:
            return
        if self.cooldown_counter > 0:
            self.cooldown_counter -= 1
            self.wait = 0
        if self._improved(current):
            self.best = current
            self.wait = 0
        elif not self.cooldown_counter > 0:
            self.wait += 1
            if self.wait >= self.patience:
                old_lr = float(self.model.optimizer.lr)
                if old_lr > self.min_lr:
                    self.model.optimizer.lr = max(old_lr * self.factor, self.min_lr)
                    self.cooldown_counter = self.cooldown
                    self.wait = 0


class EarlyStopping(Callback):
    def __init__(self, monitor="val_loss", min_delta=0, patience=0, verbose=0, mode="auto", **_):
        super(EarlyStopping, self).__init__()
        self.monitor, self.patience = monitor, patience
        self.mode = "max" if (mode == "max" or (mode == "auto" and "acc" in monitor)) else "min"
        self.min_delta = abs(min_delta) if self.mode == "max" else -abs(min_delta)
        self.stopped_epoch = 0

    def on_train_begin(self, logs=None):
        self.wait = 0
        self.stopped_epoch = 0
        self.best = -np.inf if self.mode == "max" else np.inf

    def on_epoch_end(self, epoch, logs=None):
        current = (logs or {}).get(self.monitor)
        if current is None:
            return
        better = (current - self.min_delta > self.best) if self.mode == "max" else (current - self.min_delta < self.best)
        if better:
            self.best = current
            self.wait = 0
        else:
            self.wait += 1
            if self.wait >= self.patience:
                self.stopped_epoch = epoch
                self.model.stop_training = True


class ModelCheckpoint(Callback):
    """save_best_only checkpoint kept as a device snapshot; ``restore()`` plays keras.models.load_model."""

    def __init__(self, filepath=None, monitor="val_loss", verbose=0, save_best_only=False, mode="auto", **_):
        super(ModelCheckpoint, self).__init__()
        self.filepath, self.monitor, self.save_best_only = filepath, monitor, save_best_only
        self.mode = "max" if (mode == "max" or (mode == "auto" and ("acc" in monitor or monitor.startswith("fmeasure")))) else "min"
        self.best = -np.inf if self.mode == "max" else np.inf
        self.snapshot = None
        self.saved_epoch = None

    def on_epoch_end(self, epoch, logs=None):
        current = (logs or {}).get(self.monitor)
        if self.save_best_only:
            if current is None:
                return
            better = current > self.best if self.mode == "max" else current < self.best
            if not better:
                return
            self.best = current
        self.snapshot = self.model._need_engine().snapshot(self.snapshot)
        self.saved_epoch = epoch

    def restore(self):
        if self.snapshot is None:
            raise IOError("no checkpoint was written (no epoch improved %s)" % self.monitor)
        self.model._need_engine().restore(self.snapshot)
        return self.model
