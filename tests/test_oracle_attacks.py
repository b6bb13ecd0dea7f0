"""The oracle's input gradient (inference mode) and the two L2 attacks: finite differences and invariants."""
import numpy as np

from oracle import attacks
from oracle.net import OracleNet

from tests.util import init_weights, tiny_graphs


def _net(name="pools_and_bn"):
    graphs = tiny_graphs()
    net = graphs[name if name in graphs else sorted(graphs)[0]]()
    rng = np.random.default_rng(3)
    return net, OracleNet(net.describe(), init_weights(net, rng, jitter=True))


def test_input_gradient_matches_finite_differences_in_inference_mode():
    for name in sorted(tiny_graphs()):
        net, oracle = _net(name)
        rng = np.random.default_rng(5)
        x = rng.random((3,) + tuple(net.input_shape[1:]))
        classes = net.output_shape[-1]
        t = np.eye(classes)[rng.integers(0, classes, 3)]
        g = oracle.input_gradient(x, t)

        def loss(z):
            p = oracle.forward(z, training=False)[-1]
            q = np.clip(p / p.sum(axis=1, keepdims=True), 1e-7, 1 - 1e-7)
            return float(-(t * np.log(q)).sum() / z.shape[0])

        flat = x.reshape(-1)
        for idx in rng.choice(flat.size, size=min(12, flat.size), replace=False):
            e = np.zeros_like(flat)
            e[idx] = 1e-6
            num = (loss((flat + e).reshape(x.shape)) - loss((flat - e).reshape(x.shape))) / 2e-6
            assert abs(num - g.reshape(-1)[idx]) <= 1e-5 * max(1.0, abs(num)), (name, idx, num, g.reshape(-1)[idx])


def test_attacks_respect_their_budgets():
    net, oracle = _net()
    rng = np.random.default_rng(8)
    x = rng.random((5,) + tuple(net.input_shape[1:]))
    norm = lambda a: np.sqrt((a.reshape(a.shape[0], -1) ** 2).sum(axis=1))
    adv = attacks.fgsm_minimal(oracle, x)
    # steps are 0.1 and 0.2: the third, 0.30000000000000004 in Python floats, is already > eps (as in ART); the clip
    # at 0 can only shorten a perturbation
    assert np.all(norm(adv - x) <= 0.2 + 1e-9) and np.all(norm(adv - x) > 0.05)
    adv = attacks.pgd(oracle, x, max_iter=15)
    assert np.all(norm(adv - x) <= 1.0 + 1e-9)
    assert 0.0 <= attacks.empirical_robustness(oracle, x, adv)


def test_logit_gradient_matches_finite_differences():
    for name in sorted(tiny_graphs()):
        net, oracle = _net(name)
        rng = np.random.default_rng(6)
        x = rng.random((3,) + tuple(net.input_shape[1:]))
        classes = net.output_shape[-1]
        head = rng.standard_normal((3, classes))
        g = oracle.logit_gradient(x, head)

        def f(z):
            oracle.forward(z, training=False)
            return float((head * oracle.logits).sum())

        flat = x.reshape(-1)
        for idx in rng.choice(flat.size, size=min(12, flat.size), replace=False):
            e = np.zeros_like(flat)
            e[idx] = 1e-6
            num = (f((flat + e).reshape(x.shape)) - f((flat - e).reshape(x.shape))) / 2e-6
            assert abs(num - g.reshape(-1)[idx]) <= 1e-5 * max(1.0, abs(num)), (name, idx, num, g.reshape(-1)[idx])


def test_random_sphere_fills_the_ball():
    np.random.seed(0)
    pts = attacks.random_sphere_l2(4000, 3, 2.0)
    r = np.linalg.norm(pts, axis=1)
    assert r.max() <= 2.0 + 1e-12
    # uniform in the ball: P(r <= t) = (t / R)^d
    for t in (0.8, 1.2, 1.6):
        assert abs(np.mean(r <= t) - (t / 2.0) ** 3) <= 0.03
    assert np.abs(pts.mean(axis=0)).max() <= 0.08


def test_clever_score_is_bounded_and_reproducible():
    net, oracle = _net()
    rng = np.random.default_rng(9)
    x = rng.random(tuple(net.input_shape[1:]))
    np.random.seed(42)
    s1, m1 = attacks.clever_u(oracle, x, nb_batches=4, batch_size=3, radius=2.0, pool_factor=3)
    np.random.seed(42)
    s2, m2 = attacks.clever_u(oracle, x, nb_batches=4, batch_size=3, radius=2.0, pool_factor=3)
    assert s1 == s2 and np.array_equal(m1, m2)
    assert 0.0 <= s1 <= 2.0
    assert m1.shape == (net.output_shape[-1] - 1, 4) and np.all(m1 > 0)
    # the margin over a bound on the gradient norm is a lower bound on the distance to the decision boundary:
    # the score can never exceed margin / (smallest observed maximum)
    oracle.forward(x[None], training=False)
    z = oracle.logits[0]
    pred = int(np.argmax(z))
    others = [j for j in range(z.shape[0]) if j != pred]
    assert s1 <= min(2.0, min((z[pred] - z[j]) / m1[t].min() for t, j in enumerate(others))) + 1e-12


def test_carlini_l2_keeps_the_best_adversarial_example():
    net, oracle = _net()
    rng = np.random.default_rng(8)
    x = rng.random((5,) + tuple(net.input_shape[1:]))
    labels = np.argmax(oracle.forward(x)[-1], axis=1)
    # tanh space round trip
    t = attacks.to_tanh_space(x, 0.0, 1.0)
    assert np.abs(attacks.from_tanh_space(t, 0.0, 1.0) - x).max() <= 1e-9
    # with the reference's settings (clip range 0..255 around data in [0, 1], c <= 5.12, 10 iterations of lr 0.01)
    # the tanh variable sits in its flat tail: nothing moves and the clean input comes back
    adv = attacks.carlini_l2(oracle, x)
    assert np.abs(adv - x).max() <= 1e-6
    # a stronger setting on the data's own range finds adversarial examples and returns the closest one it saw
    adv = attacks.carlini_l2(oracle, x, clip=(0.0, 1.0), initial_const=10.0, learning_rate=0.1)
    assert adv.min() >= -1e-6 and adv.max() <= 1.0 + 1e-6      # tanh / 0.999999 overshoots the range by 5e-7 of its width
    fooled = np.argmax(oracle.forward(adv)[-1], axis=1) != labels
    assert fooled.any()
    moved = np.linalg.norm((adv - x).reshape(5, -1), axis=1)
    assert np.all(moved[fooled] > 0) and np.all(moved[~fooled] <= 1e-6)
