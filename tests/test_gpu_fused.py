"""The fused step (csrc/fused_step.cu, featurenet_b200/fused.py): one persistent launch per optimisation step for
candidates made of small-channel layers.  It must compute what the separate launches compute -- checked against the
engine with FNET_FUSED=0 on identical weights, batches and permutations (training histories, every weight after several
Adam steps, evaluate), against the float64 oracle step by step, and for run-to-run determinism."""
import contextlib
import copy
import io
import os

import numpy as np
import pytest

from oracle.net import OracleNet
from featurenet_b200 import layers as L
from tests.util import init_weights

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def head(t, classes=10):
    return L.Dense(classes, activation="softmax", name="out")(L.Flatten()(t))


def cell_like():
    """What product #4 of the main_1block_nas sample is: 1x1 3->6 relu, 1x1 6->3 aligner, concat with the input."""
    x = L.Input((8, 8, 3))
    a = L.Conv2D(6, (1, 1), padding="same", activation="relu")(x)
    b = L.Conv2D(3, (1, 1), padding="same")(a)
    return L.Model(inputs=x, outputs=head(L.Concatenate(axis=1)([x, b])))


def strided_taps():
    x = L.Input((9, 7, 3))
    a = L.Conv2D(4, (2, 2), strides=(2, 2), padding="same", activation="tanh")(x)        # 13 filter rows x 4
    b = L.Conv2D(2, (1, 3), strides=(1, 1), padding="same", activation="sigmoid")(a)     # 13 filter rows x 2
    c = L.Conv2D(3, (3, 1), strides=(1, 2), padding="valid")(b)                          # 7 rows x 3, valid padding
    return L.Model(inputs=x, outputs=head(c, 7))


def head_only():
    x = L.Input((6, 5, 2))
    return L.Model(inputs=x, outputs=head(x, 4))


def nas1_product(index):
    from model.keras_model import KerasFeatureModel
    from products_tree import ProductSet
    with contextlib.redirect_stdout(io.StringIO()):
        product, original = ProductSet(os.path.join(ROOT, "featurenet_b200/data/main_1block_nas_5cells_seed0.pdt")).format_products()[index]
        return KerasFeatureModel.parse_feature_model(copy.deepcopy(product), product_features=original).build((32, 32, 3), 10)


GRAPHS = {"cell_like": cell_like, "strided_taps": strided_taps, "head_only": head_only,
          "nas1#0": lambda: nas1_product(0), "nas1#3": lambda: nas1_product(3), "nas1#4": lambda: nas1_product(4)}


def problem(rng, n, shape, classes):
    x = rng.random((n,) + shape).astype(np.float32)
    teacher = rng.standard_normal((int(np.prod(shape)), classes))
    return x, np.eye(classes, dtype=np.float32)[np.argmax((x.reshape(n, -1) - 0.5) @ teacher, axis=1)]


def run(name, fused, monkeypatch, optimizer=None):
    from featurenet_b200.optimizers import Adam
    monkeypatch.setenv("FNET_FUSED", "1" if fused else "0")
    net = GRAPHS[name]()
    shape, classes = tuple(net.input_shape[1:]), net.output_shape[-1]
    rng = np.random.default_rng(41)
    weights = init_weights(net, rng, jitter=True)
    x, y = problem(rng, 300, shape, classes)
    xv, yv = problem(rng, 150, shape, classes)
    perms = [rng.permutation(300) for _ in range(2)]
    net.compile(loss="categorical_crossentropy", metrics=["accuracy"], optimizer=optimizer or Adam(lr=1e-3))
    eng = net._need_engine()
    eng.set_weights([w.astype(np.float32) for ws in weights for w in ws])
    hist = net.fit(x, y, batch_size=128, epochs=2, validation_data=(xv, yv), permutations=perms)     # 128 + 128 + 44 rows
    score = net.evaluate(xv, yv)
    return eng, hist.history, eng.get_weights(), score


@pytest.mark.parametrize("name", sorted(GRAPHS))
def test_fused_step_equals_separate_launches(name, monkeypatch):
    eng_f, hist_f, w_f, score_f = run(name, True, monkeypatch)
    assert eng_f.fused_summary, "this graph is made of fusable layers only: the engine must have fused its step"
    assert all(eng_f.launches_per_step[key] == 1 for key in eng_f.fused_summary)
    eng_s, hist_s, w_s, score_s = run(name, False, monkeypatch)
    assert not eng_s.fused_summary and max(eng_s.launches_per_step.values()) > 1
    for key in ("loss", "val_loss"):
        assert np.allclose(hist_f[key], hist_s[key], rtol=2e-5, atol=0), key
    for key in ("acc", "val_acc"):
        assert np.abs(np.array(hist_f[key]) - np.array(hist_s[key])).max() <= 1.0 / 150 + 1e-9, key
    for a, b in zip(w_f, w_s):
        # six Adam steps on both sides; the two paths sum in different (fixed) orders
        assert np.abs(a - b).max() <= 2e-5 * np.abs(b).max() + 0.02 * 1e-3
    assert abs(score_f[0] - score_s[0]) <= 2e-5 * max(1.0, abs(score_s[0])) and abs(score_f[1] - score_s[1]) <= 1.0 / 150 + 1e-9
    # and the same run again gives the same bits
    _, hist_g, w_g, _ = run(name, True, monkeypatch)
    assert hist_g == hist_f and all(np.array_equal(a, b) for a, b in zip(w_f, w_g))


@pytest.mark.parametrize("name", ["cell_like", "strided_taps"])
def test_fused_step_matches_the_float64_oracle(name, monkeypatch):
    from featurenet_b200.optimizers import Adam
    monkeypatch.setenv("FNET_FUSED", "1")
    net = GRAPHS[name]()
    rows = net.describe()
    shape, classes = tuple(net.input_shape[1:]), net.output_shape[-1]
    rng = np.random.default_rng(3)
    weights = init_weights(net, rng, jitter=True)
    net.compile(loss="categorical_crossentropy", metrics=["accuracy"], optimizer=Adam(lr=1e-3))
    eng = net._need_engine()
    eng.set_weights([w.astype(np.float32) for ws in weights for w in ws])
    oracle = OracleNet(rows, [[w.astype(np.float32).astype(np.float64) for w in ws] for ws in weights])
    for step in range(2):
        x, t = problem(rng, 12, shape, classes)
        loss_e, acc_e = eng.train_on_batch(x, t)
        loss_o, correct_o = oracle.train_step(x.astype(np.float64), t.astype(np.float64), float(np.float32(1e-3)), {})
        assert eng.fused_summary
        assert abs(loss_e * 12 - loss_o) <= 1e-4 * max(1.0, abs(loss_o)) and round(acc_e * 12) == correct_o
        torch.cuda.synchronize()
        g = eng.grads.cpu().numpy()
        for (i, k), (off, size) in eng.w_off.items():
            want = oracle.last_grads[i][k]
            got = g[off: off + size].reshape(want.shape)
            assert np.abs(got - want).max() <= 1e-4 * max(np.abs(want).max(), 1e-12), (i, k, step)
        params, m, v = (a.cpu().numpy().astype(np.float64) for a in (eng.params, eng.adam_m, eng.adam_v))
        for (i, k), (off, size) in eng.w_off.items():
            shp = oracle.weights[i][k].shape
            oracle.weights[i][k] = params[off: off + size].reshape(shp).copy()
            oracle.adam_m[i][k] = m[off: off + size].reshape(shp).copy()
            oracle.adam_v[i][k] = v[off: off + size].reshape(shp).copy()


def test_fused_step_with_sgd_and_streamed_uint8_batches(monkeypatch):
    from featurenet_b200.optimizers import SGD
    eng_f, hist_f, w_f, _ = run("cell_like", True, monkeypatch, optimizer=SGD(lr=0.05, momentum=0.9, nesterov=True))
    eng_s, hist_s, w_s, _ = run("cell_like", False, monkeypatch, optimizer=SGD(lr=0.05, momentum=0.9, nesterov=True))
    assert eng_f.fused_summary and np.allclose(hist_f["loss"], hist_s["loss"], rtol=2e-5)
    for a, b in zip(w_f, w_s):
        assert np.abs(a - b).max() <= 5e-5 * max(np.abs(b).max(), 1e-3)

    # host batches of uint8 pixels through fit_generator: the fused kernel's first operation scales them on the device
    class Feed(object):
        scale = 255.0

        def __init__(self):
            rng = np.random.default_rng(8)
            self.px = rng.integers(0, 256, size=(4, 32, 8, 8, 3), dtype=np.uint8)
            self.labels = np.eye(10, dtype=np.float32)[rng.integers(0, 10, (4, 32))]
            self.i = 0

        def __iter__(self):
            return self

        def __next__(self):
            k = self.i % 4
            self.i += 1
            return self.px[k], self.labels[k]

    results = []
    for fused in (True, False):
        monkeypatch.setenv("FNET_FUSED", "1" if fused else "0")
        from featurenet_b200.optimizers import Adam
        net = cell_like()
        weights = init_weights(net, np.random.default_rng(5))
        net.compile(loss="categorical_crossentropy", metrics=["accuracy"], optimizer=Adam(lr=1e-3))
        net._need_engine().set_weights([w.astype(np.float32) for ws in weights for w in ws])
        hist = net.fit_generator(Feed(), steps_per_epoch=8, epochs=1)
        results.append((hist.history["loss"][0], net._need_engine().get_weights(), bool(net._need_engine().fused_summary)))
    assert results[0][2] and not results[1][2]
    assert abs(results[0][0] - results[1][0]) <= 2e-5 * abs(results[1][0])
    for a, b in zip(results[0][1], results[1][1]):
        assert np.abs(a - b).max() <= 2e-5 * np.abs(b).max() + 0.02 * 1e-3
