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
