"""Synthetic stand-ins for MNIST / CIFAR (no dataset can be downloaded on the build or GPU boxes).

Same shapes, sizes and preprocessing as /root/reference/tensorflow_generator.py:298-313: uint8
pixels -> float32 / 255, integer labels -> one-hot float32.  Labels come from a fixed random
linear teacher on the pixels, so accuracy is learnable and an accuracy band means something
(SURVEY.md section 8d).  Seeds: pixels `seed`, teacher `seed + 1`.
"""
import numpy as np

SHAPES = {"mnist": (28, 28, 1), "cifar": (32, 32, 3), "cifar10": (32, 32, 3), "cifar100": (32, 32, 3)}
SIZES = {"mnist": (60000, 10000), "cifar": (50000, 10000), "cifar10": (50000, 10000), "cifar100": (50000, 10000)}
CLASSES = {"mnist": 10, "cifar": 10, "cifar10": 10, "cifar100": 100}


def make_synth_uint8(shape, n_train, n_test, classes=10, seed=1234):
    rng = np.random.default_rng(seed)
    d = int(np.prod(shape))
    x = rng.integers(0, 256, size=(n_train + n_test, d), dtype=np.uint8)
    teacher = np.random.default_rng(seed + 1).standard_normal((d, classes)).astype(np.float32)
    labels = np.empty(n_train + n_test, np.int64)
    for lo in range(0, x.shape[0], 8192):                       # chunked: 60000 x 3072 floats would be 700 MB
        chunk = x[lo: lo + 8192].astype(np.float32) / 255.0 - 0.5
        labels[lo: lo + 8192] = np.argmax(chunk @ teacher, axis=1)
    x = x.reshape((-1,) + tuple(shape))
    return (x[:n_train], labels[:n_train]), (x[n_train:], labels[n_train:])


def to_categorical(labels, classes):
    out = np.zeros((labels.shape[0], classes), np.float32)
    out[np.arange(labels.shape[0]), labels] = 1.0
    return out


def load(dataset, n_train=None, n_test=None, seed=1234):
    """Returns (x_train, y_train), (x_test, y_test) already scaled / one-hot, like init_dataset leaves them."""
    shape, classes = SHAPES[dataset], CLASSES[dataset]
    n_train = SIZES[dataset][0] if n_train is None else n_train
    n_test = SIZES[dataset][1] if n_test is None else n_test
    (xtr, ytr), (xte, yte) = make_synth_uint8(shape, n_train, n_test, classes, seed)
    xtr = xtr.astype("float32")
    xte = xte.astype("float32")
    xtr /= 255
    xte /= 255
    return (xtr, to_categorical(ytr, classes)), (xte, to_categorical(yte, classes))
