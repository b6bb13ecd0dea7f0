"""TEST INFRASTRUCTURE ONLY -- float64 restatement of the two L2 gradient attacks behind the reference's
robustness fitness (tensorflow_generator.py:151-218 -> model/metrics.py:57-91).  PARITY UNPINNED: the
attacks themselves live in adversarial-robustness-toolbox 0.9.0 (requirements.txt:7), which is neither
vendored under /root/reference nor installable here; what follows restates its published algorithm
(FastGradientMethod with minimal=True, ProjectedGradientDescent, both with norm=2, tol=1e-7).
Nothing under featurenet_b200/ imports this module.
"""
import numpy as np

TOL = 1e-7


def _normalised(g):
    n = np.sqrt(np.sum(g.reshape(g.shape[0], -1) ** 2, axis=1)) + TOL
    return g / n.reshape((-1,) + (1,) * (g.ndim - 1))


def _predict(net, x):
    return net.forward(x, training=False)[-1]


def fgsm_minimal(net, x, eps=0.3, eps_step=0.1, clip=(0.0, 255.0)):
    x = np.asarray(x, np.float64)
    labels = np.argmax(_predict(net, x), axis=1)
    onehot = np.eye(_predict(net, x).shape[1])[labels]
    direction = _normalised(net.input_gradient(x, onehot))
    adv = x.copy()
    active = np.ones(x.shape[0], bool)
    step = eps_step
    while active.any() and step <= eps:
        adv[active] = np.clip(x[active] + step * direction[active], clip[0], clip[1])
        active = np.argmax(_predict(net, adv), axis=1) == labels
        step += eps_step
    return adv


def pgd(net, x, eps=1.0, eps_step=0.1, max_iter=100, clip=(0.0, 255.0)):
    x = np.asarray(x, np.float64)
    labels = np.argmax(_predict(net, x), axis=1)
    onehot = np.eye(_predict(net, x).shape[1])[labels]
    adv = x.copy()
    for _ in range(max_iter):
        delta = (adv - x) + eps_step * _normalised(net.input_gradient(adv, onehot))
        norm = np.sqrt(np.sum(delta.reshape(delta.shape[0], -1) ** 2, axis=1))
        shrink = np.where(norm > eps, eps / (norm + TOL), 1.0)
        adv = np.clip(x + delta * shrink.reshape((-1,) + (1,) * (x.ndim - 1)), clip[0], clip[1])
    return adv


def empirical_robustness(net, x, adv):
    x = np.asarray(x, np.float64)
    changed = np.argmax(_predict(net, adv), axis=1) != np.argmax(_predict(net, x), axis=1)
    if not changed.any():
        return 0.0
    flat = lambda a: a.reshape(a.shape[0], -1)
    return float(np.mean(np.linalg.norm(flat(adv - x), axis=1)[changed] / np.linalg.norm(flat(x), axis=1)[changed]))
