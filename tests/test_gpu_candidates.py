"""Whole-candidate parity ON THE CANDIDATES THAT ARE BENCHMARKED, at the benchmarked batch (128).

tests/test_gpu_network.py compares the engine with the NumPy oracle on tiny graphs and two templates at batch 4-6.
Here the candidates bench.py times -- product #4 of the main_1block_nas sample (the default workload), product #9 of
the nas_5_5_10 sample ("wide") and three products of the 32-candidate population (one with a separable convolution,
one with BatchNorm + concat, one with an Add merge) -- run one optimisation step at batch 128 against
oracle/torch_cpu.py in float64 on identical weights, batch and dropout masks: loss, correct count, every layer's
forward output, every weight gradient and the weights after the Adam step.

Also here: the end-of-training accuracy band of SURVEY.md 8(c) -- lenet5 on an MNIST-shaped synthetic subset,
engine loop vs the CPU port's loop on identical permutations, 3 seeds, |mean accuracy difference| <= 1.0 pp.
"""
import contextlib
import copy
import io
import os

import numpy as np
import pytest

from oracle import ops
from oracle.torch_cpu import TorchCpuNet
from tests.util import init_weights, build_template

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOL = 1e-4           # forward outputs, the loss and every weight gradient: the bar of north_star
BATCH = 128


def relerr(got, want):
    want = np.asarray(want, np.float64)
    scale = np.abs(want).max()
    return np.abs(np.asarray(got, np.float64) - want).max() / (scale if scale > 0 else 1.0)


def product_net(path, index):
    from model.keras_model import KerasFeatureModel
    from products_tree import ProductSet
    with contextlib.redirect_stdout(io.StringIO()):
        product, original = ProductSet(os.path.join(ROOT, path)).format_products()[index]
        return KerasFeatureModel.parse_feature_model(copy.deepcopy(product), product_features=original).build((32, 32, 3), 10)


def population_pick(kind_sets):
    """First buildable product of the population file whose layer kinds include every set in turn."""
    picks, taken = [], set()
    for wanted in kind_sets:
        for index in range(32):
            if index in taken:
                continue
            net = product_net("featurenet_b200/data/nas_5_5_10_pop32_seed7.pdt", index)
            if net and wanted <= {n.kind for n in net.layers} and net.count_params() < 3_000_000:
                picks.append(index)
                taken.add(index)
                break
        else:
            raise AssertionError("no population product with layers %s" % sorted(wanted))
    return picks


_PICKS = []


def pop_net(which):
    """Three DIFFERENT products of the population: one with a separable convolution, one with BatchNorm and a
    concatenation, one with an Add merge."""
    if not _PICKS:
        _PICKS.extend(population_pick([{"sepconv2d"}, {"bn", "concat"}, {"add"}]))
    return product_net("featurenet_b200/data/nas_5_5_10_pop32_seed7.pdt", _PICKS[which])


CANDIDATES = {
    "nas1#4": lambda: product_net("featurenet_b200/data/main_1block_nas_5cells_seed0.pdt", 4),
    "wide#9": lambda: product_net("tests/golden/nas_5_5_10_seed0.pdt", 9),
    "pop:sepconv": lambda: pop_net(0),
    "pop:bn+concat": lambda: pop_net(1),
    "pop:add": lambda: pop_net(2),
}


@pytest.mark.parametrize("name", sorted(CANDIDATES))
def test_benchmarked_candidate_step_matches_float64_port_at_batch_128(name):
    from featurenet_b200.optimizers import Adam
    net = CANDIDATES[name]()
    assert net, "candidate must build"
    rows = net.describe()
    rng = np.random.default_rng(77)
    weights = init_weights(net, rng, jitter=True)
    w32 = [[w.astype(np.float32).astype(np.float64) for w in ws] for ws in weights]
    net.compile(loss="categorical_crossentropy", metrics=["accuracy"], optimizer=Adam(lr=1e-3))
    eng = net._need_engine()
    eng.set_weights([w.astype(np.float32) for ws in weights for w in ws])
    port = TorchCpuNet(rows, w32, dtype=torch.float64)
    x = rng.random((BATCH, 32, 32, 3)).astype(np.float32)
    t = np.eye(10, dtype=np.float32)[rng.integers(0, 10, BATCH)]
    masks = {}
    for i, row in enumerate(rows):
        if row["class"] == "Dropout":
            n = BATCH * int(np.prod(row["out_shape"][1:]))
            masks[i] = ops.dropout_keep_mask(n, row["cfg"]["rate"], eng.dropout_seed, i, 0)
    loss_e, acc_e = eng.train_on_batch(x, t)
    torch.cuda.synchronize()
    # ReLU outputs of the engine decide which gradient paths are open on BOTH sides: an element within fp32 rounding of
    # zero that lands on the other side of the kink is a different INPUT of the backward pass (one such element among
    # 50 M moves single entries of a filter gradient by 2e-3 of its largest entry -- measured on the "wide" candidate with
    # the plain FFMA kernels); how many there are is bounded below
    relu_masks, flips = {}, 0
    for i, row in enumerate(rows):
        if row["cfg"].get("activation") == "relu":
            relu_masks[i] = (eng.act[eng._resolve(i, True)][:BATCH] > 0).cpu().numpy().reshape((BATCH,) + tuple(row["out_shape"][1:]))
    loss_o, correct_o = port.train_step(x, t, float(np.float32(1e-3)), masks, relu_masks=relu_masks)
    probe = TorchCpuNet(rows, w32, dtype=torch.float64)
    with torch.no_grad():
        plain = probe.forward(torch.as_tensor(x, dtype=torch.float64), True, masks)
    for i in relu_masks:
        y = plain[i].permute(0, 2, 3, 1).numpy() if plain[i].dim() == 4 else plain[i].numpy()
        wrong = (y > 0) != relu_masks[i]
        assert np.abs(y[wrong]).max(initial=0.0) <= 1e-5 * np.abs(y).max(), "a flipped ReLU element must be within rounding of zero"
        flips += int(wrong.sum())
    assert flips <= 64, "too many ReLU sign disagreements to be rounding: %d" % flips
    assert abs(loss_e * BATCH - loss_o) <= TOL * max(1.0, abs(loss_o))
    assert abs(round(acc_e * BATCH) - correct_o) <= 1          # an argmax tie within rounding may flip one sample
    for i, row in enumerate(rows):
        got = eng.act[eng._resolve(i, True)][:BATCH].cpu().numpy().reshape(port.out_nhwc(i).shape)
        assert relerr(got, port.out_nhwc(i)) <= TOL, "layer %d (%s) forward" % (i, row["class"])
    g = eng.grads.cpu().numpy()
    want = port.grads_keras()
    for (i, k), (off, size) in eng.w_off.items():
        got = g[off: off + size].reshape(eng.nodes[i].weights[k].shape)
        assert relerr(got, want[(i, k)]) <= TOL, "gradient of layer %d (%s) weight %d" % (i, rows[i]["class"], k)
    for a, b in zip(eng.get_weights(), port.weights_keras()):
        assert np.abs(a - b).max() <= TOL * np.abs(b).max() + 0.05 * 1e-3


def test_lenet5_end_of_training_accuracy_band():
    """BASELINE configs[0]: lenet5, MNIST shape, batch 128 -- 3 epochs on a 6 000 / 1 000 subset of the synthetic set,
    3 seeds (weights and permutation), engine vs the CPU port with the same weights and the same permutation.
    Band: |mean test accuracy difference| <= 1.0 percentage point."""
    from featurenet_b200 import synth
    from featurenet_b200.optimizers import Adam
    (x, y), (xt, yt) = synth.load("mnist", 6000, 1000)
    accs_e, accs_o = [], []
    for seed in (11, 12, 13):
        net = build_template("lenet5", (28, 28, 1))
        rng = np.random.default_rng(seed)
        weights = init_weights(net, rng)
        perm = rng.permutation(6000)
        net.compile(loss="categorical_crossentropy", metrics=["accuracy"], optimizer=Adam(lr=1e-3))
        eng = net._need_engine()
        eng.set_weights([w.astype(np.float32) for ws in weights for w in ws])
        net.fit(x, y, batch_size=BATCH, epochs=3, permutations=[perm, perm[::-1], np.roll(perm, 17)])
        accs_e.append(net.evaluate(xt, yt)[1])
        port = TorchCpuNet(net.describe(), weights)
        for order in (perm, perm[::-1], np.roll(perm, 17)):
            port.fit_epoch(x, y, BATCH, order, lr=float(np.float32(1e-3)))
        accs_o.append(port.evaluate(xt, yt)[1])
        net.release()
    assert np.mean(accs_o) > 0.15, "the task must be learnable for a band to mean anything: %r" % accs_o
    assert abs(np.mean(accs_e) - np.mean(accs_o)) <= 0.01, (accs_e, accs_o)
    assert max(abs(a - b) for a, b in zip(accs_e, accs_o)) <= 0.03, (accs_e, accs_o)


@pytest.mark.parametrize("name", sorted(CANDIDATES))
def test_no_launch_reads_what_no_launch_of_the_step_wrote(name, monkeypatch):
    """FNET_POISON=1 fills every activation, gradient and workspace buffer with NaN at allocation: a launch that reads a
    buffer no earlier launch of the step wrote (a gradient nobody produced -- round 1's dropped cell branches left the
    convolutions training on whatever the allocator handed out) turns the loss or a weight into NaN."""
    from featurenet_b200.optimizers import Adam
    monkeypatch.setenv("FNET_POISON", "1")
    net = CANDIDATES[name]()
    rng = np.random.default_rng(3)
    net.compile(loss="categorical_crossentropy", metrics=["accuracy"], optimizer=Adam(lr=1e-3))
    eng = net._need_engine()
    assert eng.poison
    x = rng.random((300, 32, 32, 3)).astype(np.float32)
    y = np.eye(10, dtype=np.float32)[rng.integers(0, 10, 300)]
    hist = net.fit(x, y, batch_size=128, epochs=2, validation_data=(x[:100], y[:100]))
    assert all(np.isfinite(v) for key in ("loss", "val_loss", "acc", "val_acc") for v in hist.history[key]), hist.history
    assert all(np.isfinite(w).all() for w in eng.get_weights())
    torch.cuda.synchronize()
    assert bool(torch.isfinite(eng.grads).all())
