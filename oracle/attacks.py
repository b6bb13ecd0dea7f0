"""TEST INFRASTRUCTURE ONLY -- float64 restatement of the two L2 gradient attacks behind the reference's
robustness fitness (tensorflow_generator.py:151-218 -> model/metrics.py:57-91).  PARITY UNPINNED: the
attacks themselves live in adversarial-robustness-toolbox 0.9.0 (requirements.txt:7), which is neither
vendored under /root/reference nor installable here; what follows restates its published algorithm
(FastGradientMethod with minimal=True, ProjectedGradientDescent, both with norm=2, tol=1e-7).  The CLEVER score at
the end follows the reference's OWN code (model/metrics.py:205-324); only its sampler comes from ART.
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


# ---------------------------------------------------------------------------------------------- CLEVER
# model/metrics.py:205-324 (clever_u -> clever_t), which IS under /root/reference; the two pieces it takes from
# ART 0.9.0 are `random_sphere` (restated below from the published code) and `class_gradient(logits=True)`
# (OracleNet.logit_gradient).  Random draws come from the GLOBAL numpy stream in the reference's order: per target
# class one randn(pool, dim), then nb_batches draws of np.random.choice(pool, batch_size).

def random_sphere_l2(nb_points, nb_dims, radius):
    """ART 0.9.0 art/utils.py random_sphere(norm=2): Gaussian directions, radii from the regularised lower
    incomplete gamma function so that the points fill the ball uniformly."""
    from scipy.special import gammainc
    a = np.random.randn(nb_points, nb_dims)
    s2 = np.sum(a ** 2, axis=1)
    base = gammainc(nb_dims / 2.0, s2 / 2.0) ** (1.0 / nb_dims) * radius / np.sqrt(s2)
    return a * base[:, None]


def weibull_location(grad_norms, c_init=1.0):
    """model/metrics.py:316-317: location of a reverse-Weibull MLE fit to the per-batch maxima."""
    from scipy.optimize import fmin
    from scipy.stats import weibull_min
    _, loc, _ = weibull_min.fit(-np.asarray(grad_norms, np.float64), c_init, optimizer=fmin)
    return loc


def clever_t(net, x, target, nb_batches, batch_size, radius, c_init=1.0, pool_factor=10, clip=(0.0, 255.0)):
    """model/metrics.py:242-324 for norm=2 (its dual is 2).  Returns (score, per-batch maxima of the gradient norm)."""
    x = np.asarray(x, np.float64)
    net.forward(x[None], training=False)
    z = net.logits[0]
    pred = int(np.argmax(z))
    if target == pred:
        raise ValueError("The targeted class is the predicted class.")
    pool_n = pool_factor * batch_size
    pool = random_sphere_l2(pool_n, x.size, radius).reshape((pool_n,) + x.shape) + x[None]
    pool = np.clip(pool.astype(np.float32), clip[0], clip[1]).astype(np.float64)
    head = np.zeros((batch_size, z.shape[0]))
    head[:, pred], head[:, target] = 1.0, -1.0
    maxima = []
    for _ in range(nb_batches):
        xs = pool[np.random.choice(pool_n, batch_size)]
        g = net.logit_gradient(xs, head).reshape(batch_size, -1)
        maxima.append(np.max(np.linalg.norm(g, axis=1)))
    loc = weibull_location(maxima, c_init)
    value = z[pred] - z[target]
    return min(-value / loc, radius), np.asarray(maxima)


def clever_u(net, x, nb_batches, batch_size, radius, c_init=1.0, pool_factor=10, clip=(0.0, 255.0)):
    """model/metrics.py:205-239: the smallest targeted score over the classes other than the predicted one."""
    x = np.asarray(x, np.float64)
    net.forward(x[None], training=False)
    pred = int(np.argmax(net.logits[0]))
    scores, maxima = [], []
    for j in range(net.logits.shape[1]):
        if j != pred:
            s, m = clever_t(net, x, j, nb_batches, batch_size, radius, c_init, pool_factor, clip)
            scores.append(s)
            maxima.append(m)
    return float(np.min(scores)), np.asarray(maxima)


# ---------------------------------------------------------------------------------------------- Carlini-Wagner L2
# ART 0.9.0 CarliniL2Method (un-vendored: restated from the published algorithm, PARITY UNPINNED), with the settings
# the reference reaches it with (model/metrics.py:45-55, tensorflow_generator.py:155-167): untargeted, labels = the
# model's predictions, confidence 0, learning_rate 0.01, binary_search_steps 10, max_iter 10, initial_const 0.01,
# max_halving 5, max_doubling 5, clip_values (0, 255).  `norm` and `minimal` are not parameters of this attack
# and set_params drops them.

C_UPPER_BOUND = 10e10
TANH_SMOOTHER = 0.999999


def to_tanh_space(x, lo, hi):
    u = (np.clip(x, lo, hi) - lo) / (hi - lo)
    return np.arctanh((u * 2 - 1) * TANH_SMOOTHER)


def from_tanh_space(t, lo, hi):
    return (np.tanh(t) / TANH_SMOOTHER + 1) / 2 * (hi - lo) + lo


def _cw_logits(net, x):
    net.forward(x, training=False)
    return net.logits.copy()


def _cw_loss(net, x, x_adv, target, c, confidence):
    l2 = np.sum(np.square(x - x_adv).reshape(x.shape[0], -1), axis=1)
    z = _cw_logits(net, x_adv.astype(np.float32).astype(np.float64))
    z_target = np.sum(z * target, axis=1)
    z_other = np.max(z * (1 - target) + (np.min(z, axis=1) - 1)[:, None] * target, axis=1)
    hinge = np.maximum(z_target - z_other + confidence, 0.0)          # untargeted
    return z, l2, c * hinge + l2


def _cw_direction(net, z, target, x, x_adv, x_adv_tanh, c, lo, hi):
    """minus the gradient of the loss with respect to the tanh-space variable"""
    i_add = np.argmax(target, axis=1)
    i_sub = np.argmax(z * (1 - target) + (np.min(z, axis=1) - 1)[:, None] * target, axis=1)
    head = np.zeros_like(z)
    head[np.arange(z.shape[0]), i_add] += 1.0
    head[np.arange(z.shape[0]), i_sub] -= 1.0
    g = net.logit_gradient(x_adv, head)
    g = g * c.reshape((-1,) + (1,) * (x.ndim - 1))
    g = g + 2 * (x_adv - x)
    g = g * (hi - lo)
    g = g * (1 - np.square(np.tanh(x_adv_tanh))) / (2 * TANH_SMOOTHER)
    return -g


def carlini_l2(net, x, confidence=0.0, learning_rate=0.01, binary_search_steps=10, max_iter=10, initial_const=0.01,
               max_halving=5, max_doubling=5, clip=(0.0, 255.0)):
    x = np.asarray(x, np.float32).astype(np.float64)
    lo, hi = clip
    n = x.shape[0]
    wide = lambda v: v.reshape((-1,) + (1,) * (x.ndim - 1))
    probs = net.forward(x, training=False)[-1]
    y = np.eye(probs.shape[1])[np.argmax(probs, axis=1)]
    x_tanh = to_tanh_space(x, lo, hi)
    c = initial_const * np.ones(n)
    c_lower = np.zeros(n)
    c_double = np.ones(n, bool)
    best_l2 = np.inf * np.ones(n)
    best_adv = x.copy()
    for _ in range(binary_search_steps):
        if int(np.sum(c < C_UPPER_BOUND)) == 0:
            break
        lr = learning_rate * np.ones(n)
        adv, adv_tanh = x.copy(), x_tanh.copy()
        z, l2, loss = _cw_loss(net, x, adv, y, c, confidence)
        success = loss - l2 <= 0
        overall = success
        for _ in range(max_iter):
            better = success & (l2 < best_l2)
            best_l2[better] = l2[better]
            best_adv[better] = adv[better]
            active = (c < C_UPPER_BOUND) & (lr > 0)
            if int(np.sum(active)) == 0:
                break
            step = _cw_direction(net, z[active], y[active], x[active], adv[active], adv_tanh[active], c[active], lo, hi)
            prev_loss, best_loss = loss.copy(), loss.copy()
            best_lr = np.zeros(n)
            halving = np.zeros(n)
            for _ in range(max_halving):
                need = loss[active] >= prev_loss[active]
                if int(np.sum(need)) == 0:
                    break
                sel = active.copy()
                sel[active] = need
                trial = from_tanh_space(adv_tanh[sel] + wide(lr[sel]) * step[need], lo, hi)
                _, l2[sel], loss[sel] = _cw_loss(net, x[sel], trial, y[sel], c[sel], confidence)
                lower = loss < best_loss
                best_lr[lower] = lr[lower]
                best_loss[lower] = loss[lower]
                lr[sel] /= 2
                halving[sel] += 1
            lr[active] *= 2
            for _ in range(max_doubling):
                need = (halving[active] == 1) & (loss[active] <= best_loss[active])
                if int(np.sum(need)) == 0:
                    break
                sel = active.copy()
                sel[active] = need
                lr[sel] *= 2
                trial = from_tanh_space(adv_tanh[sel] + wide(lr[sel]) * step[need], lo, hi)
                _, l2[sel], loss[sel] = _cw_loss(net, x[sel], trial, y[sel], c[sel], confidence)
                lower = loss < best_loss
                best_lr[lower] = lr[lower]
                best_loss[lower] = loss[lower]
            lr[halving == 1] /= 2
            move = best_lr[active] > 0
            if int(np.sum(move)) > 0:
                sel = active.copy()
                sel[active] = move
                adv_tanh[sel] = adv_tanh[sel] + wide(best_lr[sel]) * step[move]
                adv[sel] = from_tanh_space(adv_tanh[sel], lo, hi)
                z[sel], l2[sel], loss[sel] = _cw_loss(net, x[sel], adv[sel], y[sel], c[sel], confidence)
                success = loss - l2 <= 0
                overall = overall | success
        better = success & (l2 < best_l2)
        best_l2[better] = l2[better]
        best_adv[better] = adv[better]
        c_double[overall] = False
        c[overall] = (c_lower + c)[overall] / 2
        c_old = c
        c[~overall & c_double] *= 2
        c[~overall & ~c_double] += (c - c_lower)[~overall & ~c_double] / 2
        c_lower[~overall] = c_old[~overall]
    return best_adv
