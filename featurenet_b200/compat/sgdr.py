"""Cosine-annealing learning-rate schedule with warm restarts (SGDR), as a callback.

Behaviour of /root/reference/sgdr.py:5-86 (defined there, imported by nothing): per batch
``lr = min + (max - min)/2 * (1 + cos(pi * batches_since_restart / (steps_per_epoch * cycle_length)))``;
when ``epoch + 1 == next_restart`` the cycle restarts, grows by ``mult_factor``, ``max_lr`` decays
and the weights are remembered; training ends on the weights of the last completed cycle.
The schedule is host arithmetic; the weight memory is a device-to-device snapshot.
"""
import numpy as np

from featurenet_b200.callbacks import Callback


class SGDRScheduler(Callback):
    def __init__(self, min_lr, max_lr, steps_per_epoch, lr_decay=1, cycle_length=10, mult_factor=2):
        super(SGDRScheduler, self).__init__()
        self.min_lr, self.max_lr, self.lr_decay = min_lr, max_lr, lr_decay
        self.batch_since_restart = 0
        self.next_restart = cycle_length
        self.steps_per_epoch = steps_per_epoch
        self.cycle_length, self.mult_factor = cycle_length, mult_factor
        self.history = {}
        self.best_weights = None

    def clr(self):
        progress = self.batch_since_restart / (self.steps_per_epoch * self.cycle_length)
        return self.min_lr + 0.5 * (self.max_lr - self.min_lr) * (1 + np.cos(progress * np.pi))

    def on_train_begin(self, logs=None):
        self.model.optimizer.lr = float(self.max_lr)

    def on_batch_end(self, batch, logs=None):
        self.history.setdefault("lr", []).append(float(self.model.optimizer.lr))
        for k, v in (logs or {}).items():
            self.history.setdefault(k, []).append(v)
        self.batch_since_restart += 1
        self.model.optimizer.lr = float(self.clr())

    def on_epoch_end(self, epoch, logs=None):
        if epoch + 1 == self.next_restart:
            self.batch_since_restart = 0
            self.cycle_length = np.ceil(self.cycle_length * self.mult_factor)
            self.next_restart += self.cycle_length
            self.max_lr *= self.lr_decay
            self.best_weights = self.model._need_engine().snapshot(self.best_weights)

    def on_train_end(self, logs=None):
        # the reference raises AttributeError when no cycle ever completed; same here
        if self.best_weights is None:
            raise AttributeError("'SGDRScheduler' object has no attribute 'best_weights'")
        self.model._need_engine().restore(self.best_weights)
