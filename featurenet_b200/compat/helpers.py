"""``train_model``: the training loop every candidate goes through.

Behaviour of /root/reference/helpers.py:45-175 on the CUDA engine: callbacks
[ReduceLROnPlateau, LearningRateScheduler(lr_schedule), EarlyStopping(val_acc)] when
``scheduler`` is set, plus a best-val_loss checkpoint that is restored after ``fit``
(the reference writes and reloads saved_models/temp_model.h5; here it is a device snapshot).
An out-of-memory candidate yields ``(None, history)`` instead of raising.
"""
import math
import os
import threading

import numpy as np
import torch

from featurenet_b200.callbacks import (Callback, EarlyStopping, LearningRateScheduler, ModelCheckpoint,  # noqa: F401
                                       ReduceLROnPlateau)


# The reference swallows an out-of-memory candidate (helpers.py:172-174: print, model = None).  The convention is kept;
# the reason is also left here, per host thread, for a scheduler that trains several candidates side by side and has
# to tell "does not fit this GPU" from "did not fit beside its neighbours" (featurenet_b200/population.py).
_failure = threading.local()


def clear_failure():
    _failure.text = None


def last_failure():
    return getattr(_failure, "text", None)


class TimedStopping(Callback):
    """Stop when an epoch or the whole run exceeded a time budget (defined, unused: helpers.py:12-42,81)."""

    def __init__(self, generator, epoch_seconds=None, total_seconds=None, verbose=0):
        super(TimedStopping, self).__init__()
        import time
        self._clock = time.time
        self.start_time = 0
        self.epoch_seconds, self.total_seconds = epoch_seconds, total_seconds
        self.verbose, self.generator = verbose, generator

    def on_train_begin(self, logs=None):
        self.start_time = self._clock()

    def on_epoch_begin(self, epoch, logs=None):
        self.start_epoch = self._clock()

    def on_epoch_end(self, epoch, logs=None):
        over_total = self.total_seconds and self._clock() - self.start_time > self.total_seconds
        over_epoch = self.epoch_seconds and self._clock() - self.start_epoch > self.epoch_seconds
        if over_total or over_epoch:
            self.generator.stop_training = self.model.stop_training = True
            if self.verbose:
                print("Stopping after the time budget was reached.")


def lr_schedule(epoch):
    """1e-3, reduced after 80, 120, 160 and 180 epochs (helpers.py:45-67)."""
    lr = 1e-3
    if epoch > 180:
        lr *= 0.5e-3
    elif epoch > 160:
        lr *= 1e-3
    elif epoch > 120:
        lr *= 1e-2
    elif epoch > 80:
        lr *= 1e-1
    print("Learning rate: ", lr)
    return lr


def train_model(model, x_train, y_train, x_test, y_test, epochs, batch_size, scheduler=False,
                data_augmentation=False, model_path=None):
    callbacks = []
    if scheduler:
        callbacks = [ReduceLROnPlateau(factor=np.sqrt(0.1), cooldown=0, patience=5, min_lr=0.5e-6),
                     LearningRateScheduler(lr_schedule),
                     EarlyStopping(monitor="val_acc", mode="max", min_delta=0.01, patience=20)]
    checkpoint = ModelCheckpoint(model_path, monitor="val_loss", mode="min", save_best_only=True)
    callbacks.append(checkpoint)
    history = None
    try:
        if not data_augmentation:
            print("Not using data augmentation.")
            history = model.fit(x_train, y_train, batch_size=batch_size, epochs=epochs,
                                validation_data=(x_test, y_test), shuffle=True, callbacks=callbacks, verbose=2)
        else:
            print("Using real-time data augmentation.")
            from featurenet_b200.augment import ShiftFlipGenerator
            flow = ShiftFlipGenerator(x_train, y_train, batch_size=batch_size, width_shift_range=0.1,
                                      height_shift_range=0.1, horizontal_flip=True)
            history = model.fit_generator(flow, steps_per_epoch=math.ceil(len(x_train) / batch_size),
                                          validation_data=(x_test, y_test), epochs=epochs, verbose=2, workers=4,
                                          callbacks=callbacks)
        model = checkpoint.restore()          # keras.models.load_model(best checkpoint); os.remove(...)
    except torch.cuda.OutOfMemoryError as e:
        print("model exhausted memory while training, error {}".format(e))
        _failure.text = "OutOfMemoryError: {}".format(str(e)[:300])
        model = None
    return model, history
