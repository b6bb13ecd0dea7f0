"""Real-time data augmentation of helpers.train_model(data_augmentation=True), done on the device.

/root/reference/helpers.py:115-167 builds ``ImageDataGenerator(width_shift_range=0.1,
height_shift_range=0.1, horizontal_flip=True, fill_mode='nearest')`` and trains with
``fit_generator(datagen.flow(x, y, batch_size), workers=4)``: every image is shifted and flipped
on the host with scipy.  Here the random draws are made on the host once per epoch (three small
arrays) and the transform itself is fused into the batch gather kernel (csrc/augment.cu).

Keras semantics restated (keras-preprocessing 1.0.x, third-party): per image
``tx ~ U(-hr, hr) * H`` (rows), ``ty ~ U(-wr, wr) * W`` (columns), flip with probability 1/2;
the image is resampled at (row + tx, col + ty) with order-1 interpolation and edge clamping,
then flipped along the width.
"""
import numpy as np


class ShiftFlipGenerator(object):
    def __init__(self, x, y, batch_size=32, width_shift_range=0.0, height_shift_range=0.0, horizontal_flip=False,
                 seed=None):
        self.x, self.y, self.batch_size = x, y, int(batch_size)
        self.width_shift_range = float(width_shift_range)
        self.height_shift_range = float(height_shift_range)
        self.horizontal_flip = bool(horizontal_flip)
        self.rng = np.random.default_rng(seed)

    def draw(self, n):
        """Per-position (tx, ty, flip) for one epoch of n samples."""
        h, w = self.x.shape[1], self.x.shape[2]
        hr, wr = self.height_shift_range, self.width_shift_range
        tx = self.rng.uniform(-hr, hr, n) * (h if hr < 1 else 1.0) if hr else np.zeros(n)
        ty = self.rng.uniform(-wr, wr, n) * (w if wr < 1 else 1.0) if wr else np.zeros(n)
        flip = (self.rng.random(n) < 0.5) if self.horizontal_flip else np.zeros(n, bool)
        return tx.astype(np.float32), ty.astype(np.float32), flip.astype(np.uint8)
