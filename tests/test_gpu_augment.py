"""On-device shift/flip augmentation vs scipy's affine_transform (what Keras' ImageDataGenerator calls)
and the data_augmentation=True path of helpers.train_model end to end."""
import contextlib
import ctypes as C
import io

import numpy as np
import pytest
from scipy import ndimage

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def keras_shift_flip(img, tx, ty, flip):
    """keras_preprocessing.image.apply_affine_transform for a pure translation, then flip_axis(width)."""
    out = np.stack([ndimage.affine_transform(img[..., c], np.eye(2), offset=(tx, ty), order=1, mode="nearest")
                    for c in range(img.shape[-1])], axis=-1)
    return out[:, ::-1, :] if flip else out


@pytest.mark.parametrize("shape", [(28, 28, 1), (32, 32, 3), (9, 13, 5)])
def test_gather_shift_flip_matches_scipy(shape):
    from featurenet_b200 import lib as fl
    L = fl.get()
    rng = np.random.default_rng(shape[0])
    n, batch, cursor = 40, 16, 7
    src = rng.random((n,) + shape).astype(np.float32)
    perm = rng.permutation(n).astype(np.int32)
    tx = (rng.uniform(-0.1, 0.1, n) * shape[0]).astype(np.float32)
    ty = (rng.uniform(-0.1, 0.1, n) * shape[1]).astype(np.float32)
    tx[cursor], ty[cursor] = 0.0, 0.0                        # identity transform must copy exactly
    tx[cursor + 1], ty[cursor + 1] = 2.0, -3.0               # integer shift: pure edge-clamped copy
    flip = (rng.random(n) < 0.5).astype(np.uint8)
    flip[cursor] = 0
    state = torch.tensor([0, cursor, 0, 0, 0, 0, 0, 0], dtype=torch.int32, device="cuda")
    d = lambda a: torch.from_numpy(a).cuda()
    srcd, permd, txd, tyd, flipd = d(src), d(perm), d(tx), d(ty), d(flip)
    dst = torch.empty((batch,) + shape, device="cuda")
    p = lambda t: C.c_void_p(t.data_ptr())
    L.fnet_gather_shift_flip(batch, shape[0], shape[1], shape[2], p(permd), p(state), p(txd), p(tyd), p(flipd),
                             p(srcd), p(dst), None)
    torch.cuda.synchronize()
    got = dst.cpu().numpy()
    for i in range(batch):
        pos = cursor + i
        want = keras_shift_flip(src[perm[pos]].astype(np.float64), float(tx[pos]), float(ty[pos]), bool(flip[pos]))
        assert np.abs(got[i] - want).max() <= 2e-6, "sample %d" % i
    assert np.array_equal(got[0], src[perm[cursor]])


def test_train_model_with_data_augmentation_runs_and_learns(monkeypatch):
    from helpers import train_model
    from featurenet_b200.optimizers import Adam
    from tests.util import build_template
    rng = np.random.default_rng(1)
    x = rng.random((512, 28, 28, 1)).astype(np.float32)
    # labels depend on the mean brightness of the image halves: invariant enough under small shifts and flips
    y = np.eye(10, dtype=np.float32)[(x[:, :14].mean(axis=(1, 2, 3)) > x[:, 14:].mean(axis=(1, 2, 3))).astype(int)]
    net = build_template("lenet5", (28, 28, 1))
    net.compile(loss="categorical_crossentropy", metrics=["accuracy"], optimizer=Adam(lr=1e-3))
    with contextlib.redirect_stdout(io.StringIO()):
        model, hist = train_model(net, x, y, x[:128], y[:128], 3, 64, scheduler=True, data_augmentation=True)
    assert model is net and len(hist.history["loss"]) == 3
    assert ("train_aug", 64) in net._need_engine().programs           # the augmenting gather was the one used
    assert np.isfinite(hist.history["loss"]).all()
