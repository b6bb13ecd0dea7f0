"""Empirical robustness of a trained candidate under L2 gradient attacks.

The reference scores candidates with ``model/metrics.py:57-91`` (``empirical_robustness``) on top of the
attacks of adversarial-robustness-toolbox **0.9.0** (``requirements.txt:7``; un-vendored, not installable
here), called from ``tensorflow_generator.py:151-218`` with ``norm=2`` and ``clip_values=(0, 255)``.
What is restated here, from ART's published algorithm (parity with ART itself is unpinned):

* ``fgsm``: ``FastGradientMethod(norm=2, eps=0.3, eps_step=0.1, minimal=True)`` -- one loss gradient per batch
  at the clean input (labels = the model's own predictions), then ``x + k*eps_step * g/(||g||_2 + tol)`` for
  k = 1, 2, ... while the step is <= eps, each time only for the rows that are still classified as before.
* ``pgd``: ``ProjectedGradientDescent(norm=2, eps=1.0, eps_step=0.1, max_iter=100)`` -- repeated
  normalised-gradient steps from the clean input, each followed by a projection onto the L2 ball of
  radius eps around it.
* the score: mean over the rows whose prediction changed of ``||x_adv - x||_2 / ||x||_2``.

The gradient is dL/dx of the INFERENCE graph (``Engine.input_gradient``); the update, projection and clip
run in one kernel per step (``fnet_l2_attack_step``).  ``clever`` (at the end of this file) follows the
reference's own ``model/metrics.py:205-324``, and ``cw`` (Carlini-Wagner L2, after it) ART's ``CarliniL2Method``.
"""
import ctypes as C

import numpy as np
import torch

FGSM_DEFAULTS = {"eps": 0.3, "eps_step": 0.1}                       # ART 0.9.0 FastGradientMethod
PGD_DEFAULTS = {"eps": 1.0, "eps_step": 0.1, "max_iter": 100}       # tensorflow_generator.py:158-160 + ART default
CLIP_VALUES = (0.0, 255.0)                                          # tensorflow_generator.py:167
BATCH = 128                                                         # ART's default attack batch size


def _one_hot_predictions(net, x):
    p = net.predict(x)
    labels = np.argmax(p, axis=1)
    return labels, np.eye(p.shape[1], dtype=np.float32)[labels]


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def craft(net, x, attack, params=None):
    """Adversarial examples for the host array ``x`` ([n, ...] float32); returns a host array of the same shape."""
    eng = net._need_engine()
    L = eng.lib
    x = np.ascontiguousarray(x, dtype=np.float32)
    n = x.shape[0]
    row = int(np.prod(x.shape[1:]))
    cfg = dict(FGSM_DEFAULTS if attack == "fgsm" else PGD_DEFAULTS)
    cfg.update({k: v for k, v in (params or {}).items() if k in cfg})
    if attack == "cw":
        return carlini_l2(net, x, params)
    if attack not in ("fgsm", "pgd"):
        raise NotImplementedError("attack %r is not built (fgsm, pgd and cw are)" % attack)
    if (params or {}).get("norm", 2) != 2:
        raise NotImplementedError("only the L2 attacks the reference uses are built")
    adv = np.empty_like(x)
    for lo in range(0, n, BATCH):
        xb = x[lo: lo + BATCH]
        b = xb.shape[0]
        labels, onehot = _one_hot_predictions(net, xb)
        with torch.cuda.device(eng.device):
            x0 = torch.from_numpy(xb).to(eng.device)
            cur = x0.clone()
            active = torch.ones(b, dtype=torch.uint8, device=eng.device)
        if attack == "fgsm":
            grad = eng.input_gradient(x0, onehot)
            step = cfg["eps_step"]
            while bool(active.any().item()) and step <= cfg["eps"]:        # Python floats, as ART accumulates them
                L.fnet_l2_attack_step(b, row, _ptr(x0), _ptr(grad), _ptr(x0), _ptr(cur), step, 0.0, CLIP_VALUES[0],
                                      CLIP_VALUES[1], _ptr(active), eng.sp)
                eng.stream.synchronize()
                still = np.argmax(net.predict(cur.cpu().numpy()), axis=1) == labels
                active = torch.from_numpy(still.astype(np.uint8)).to(eng.device)
                step += cfg["eps_step"]
        else:
            for _ in range(int(cfg["max_iter"])):
                grad = eng.input_gradient(cur, onehot)
                L.fnet_l2_attack_step(b, row, _ptr(x0), _ptr(grad), _ptr(cur), _ptr(cur), cfg["eps_step"], cfg["eps"],
                                      CLIP_VALUES[0], CLIP_VALUES[1], None, eng.sp)
            eng.stream.synchronize()
        adv[lo: lo + b] = cur.cpu().numpy()
    return adv


def empirical_robustness(net, x, attack, params=None):
    """model/metrics.py:57-91: (mean relative L2 perturbation over the rows whose prediction changed, x_adv)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    adv = craft(net, x, attack, params)
    changed = np.argmax(net.predict(adv), axis=1) != np.argmax(net.predict(x), axis=1)
    if not changed.any():
        return 0.0, adv
    flat = lambda a: a.reshape(a.shape[0], -1).astype(np.float64)
    pert = np.linalg.norm(flat(adv - x), axis=1)[changed]
    return float(np.mean(pert / np.linalg.norm(flat(x), axis=1)[changed])), adv


def attack_robustness(net, attack, x, y, params=None):
    """tensorflow_generator.py:151-173: (empirical robustness, clean accuracy, adversarial accuracy)."""
    value, adv = empirical_robustness(net, x, attack, params)
    clean = net.evaluate(x, y, verbose=0)
    fooled = net.evaluate(adv, y, verbose=0)
    return float(value), float(clean[1]), float(fooled[1])


# ---------------------------------------------------------------------------------------------- CLEVER
# model/metrics.py:205-324 (clever_u / clever_t) as called from tensorflow_generator.py:180-203 with norm=2,
# radius 2, nb_batches 10, batch_size 5, pool_factor 3.  The reference asks ART for the full class-gradient
# tensor of every batch of 5 samples, once per target class; here the difference grads[:, pred] - grads[:, target]
# is ONE backward pass with e_pred - e_target at the logits (Engine.logit_gradient), and since every batch is
# drawn from a pool of pool_factor*batch_size points per class, the whole score of an image costs a single
# forward+backward over (classes - 1) * pool rows.  The random draws are taken from numpy's global stream in
# the reference's order (per target class: one randn for the pool, then nb_batches choices), so a caller who
# seeds numpy gets the samples the reference would have drawn.

def random_sphere_l2(nb_points, nb_dims, radius):
    """Uniform samples from the L2 ball (ART 0.9.0 ``art.utils.random_sphere`` with norm=2, restated)."""
    from scipy.special import gammainc
    a = np.random.randn(nb_points, nb_dims)
    s2 = np.sum(a ** 2, axis=1)
    return a * (gammainc(nb_dims / 2.0, s2 / 2.0) ** (1.0 / nb_dims) * radius / np.sqrt(s2))[:, None]


def _clever_maxima(net, x, nb_batches, batch_size, radius, norm, pool_factor, only_target=None):
    """The GPU half of clever_u: (margins z[pred] - z[target] per target class, per-batch gradient-norm maxima per
    target class).  All random draws of the sample happen here, in the reference's order.  only_target restricts
    the work to one target class (clever_t)."""
    if norm != 2:
        raise NotImplementedError("only norm=2, the one the reference uses (tensorflow_generator.py:186), is built")
    if pool_factor < 1:
        raise ValueError("The `pool_factor` must be larger than 1.")
    eng = net._need_engine()
    x = np.ascontiguousarray(x, dtype=np.float32)
    classes, pool_n = eng.n_out, int(pool_factor) * int(batch_size)
    zero = np.zeros((1, classes), np.float32)
    _, z = eng.logit_gradient(x[None], zero)
    z = z[0].cpu().numpy().astype(np.float64)
    pred = int(np.argmax(z))
    targets = [j for j in range(classes) if j != pred]
    if only_target is not None:
        if int(only_target) == pred:
            raise ValueError("The targeted class is the predicted class.")
        targets = [int(only_target)]
    rows = np.empty((len(targets) * pool_n,) + x.shape, np.float32)
    head = np.zeros((len(targets) * pool_n, classes), np.float32)
    picks = []
    for t, j in enumerate(targets):
        pool = random_sphere_l2(pool_n, x.size, radius).reshape((pool_n,) + x.shape) + x[None].astype(np.float64)
        rows[t * pool_n: (t + 1) * pool_n] = np.clip(pool.astype(np.float32), CLIP_VALUES[0], CLIP_VALUES[1])
        head[t * pool_n: (t + 1) * pool_n, pred] = 1.0
        head[t * pool_n: (t + 1) * pool_n, j] = -1.0
        picks.append([np.random.choice(pool_n, batch_size) for _ in range(nb_batches)])
    grad, _ = eng.logit_gradient(rows, head)
    norms = np.linalg.norm(grad.reshape(grad.shape[0], -1).cpu().numpy().astype(np.float64), axis=1)
    if np.isnan(norms).any():
        raise Exception("The classifier results NaN gradients.")
    maxima = [[float(np.max(norms[t * pool_n + idx])) for idx in picks[t]] for t in range(len(targets))]
    return [z[pred] - z[j] for j in targets], maxima


def _clever_from_fits(margins, locs, radius):
    """model/metrics.py:319-324 and :239: min over the target classes of min(-margin / loc, radius)."""
    return float(np.min([min(-m / loc, radius) for m, loc in zip(margins, locs)]))


def clever_u(net, x, nb_batches, batch_size, radius, norm=2, c_init=1, pool_factor=10, return_maxima=False):
    """Untargeted CLEVER score of ONE sample ``x`` (model/metrics.py:205-239)."""
    from . import weibull
    margins, maxima = _clever_maxima(net, x, nb_batches, batch_size, radius, norm, pool_factor)
    score = _clever_from_fits(margins, [weibull.location(m, c_init) for m in maxima], radius)
    return (score, np.asarray(maxima)) if return_maxima else score


def clever_t(net, x, target_class, nb_batches, batch_size, radius, norm=2, c_init=1, pool_factor=10):
    """Targeted CLEVER score of ONE sample for ONE target class (model/metrics.py:242-324)."""
    from . import weibull
    margins, maxima = _clever_maxima(net, x, nb_batches, batch_size, radius, norm, pool_factor, only_target=target_class)
    return _clever_from_fits(margins, [weibull.location(maxima[0], c_init)], radius)


def clever_score(net, x_set, nb_batches=10, batch_size=5, radius=2, norm=2, pool_factor=3, workers=None):
    """tensorflow_generator.py:188-203: the average untargeted CLEVER score over a set of samples.  The gradients of
    all samples first (GPU; the random draws stay in the reference's order), then every Weibull fit of the set over
    worker processes (featurenet_b200/weibull.py): same numbers as sample by sample, a fraction of the wall time."""
    from . import weibull
    per_sample = [_clever_maxima(net, element, nb_batches, batch_size, radius, norm, pool_factor) for element in x_set]
    if not per_sample:
        return float("nan")
    locs = weibull.locations([m for _, maxima in per_sample for m in maxima], 1, workers)
    scores, at = [], 0
    for margins, maxima in per_sample:
        scores.append(_clever_from_fits(margins, locs[at: at + len(maxima)], radius))
        at += len(maxima)
    return float(np.average(scores))


# ---------------------------------------------------------------------------------------------- Carlini-Wagner L2
# ART 0.9.0 ``CarliniL2Method`` as the reference reaches it (model/metrics.py:45-55 builds it with defaults, then
# tensorflow_generator.py:155-156 sets targeted=False; ``norm`` and ``minimal`` are not parameters of this attack
# and are dropped by set_params): labels = the model's own predictions, optimisation in tanh space over
# clip_values, per-sample binary search of the constant c, and per iteration a line search on the learning rate
# (halve while the loss does not drop, else double while it keeps dropping).  Restated from the published
# algorithm; ART is un-vendored, so parity with ART itself is unpinned.  Every loss evaluation is one captured
# forward of the engine (Engine.logits); the gradient is one forward+backward with e_label - e_runner_up at the
# logits (Engine.logit_gradient) instead of ART's two class_gradient calls.

CW_DEFAULTS = {"confidence": 0.0, "learning_rate": 0.01, "binary_search_steps": 10, "max_iter": 10, "initial_const": 0.01,
               "max_halving": 5, "max_doubling": 5, "batch_size": 128}
_C_LIMIT = 10e10            # binary search gives up on a sample beyond this constant
_SMOOTH = 0.999999          # keeps arctanh finite at the ends of the clip range


class _CarliniBatch(object):
    """State of the attack on one batch; every array is indexed by the row in the batch."""

    def __init__(self, net, x, cfg):
        self.eng, self.cfg = net._need_engine(), cfg
        self.lo, self.hi = CLIP_VALUES
        self.x = x.astype(np.float64)
        self.n = x.shape[0]
        labels = np.argmax(net.predict(x), axis=1)
        self.onehot = np.eye(self.eng.n_out)[labels]
        unit = (np.clip(self.x, self.lo, self.hi) - self.lo) / (self.hi - self.lo)
        self.x_tanh = np.arctanh((unit * 2 - 1) * _SMOOTH)
        self.best_l2 = np.full(self.n, np.inf)
        self.best_adv = self.x.copy()

    def image(self, t):
        return (np.tanh(t) / _SMOOTH + 1) / 2 * (self.hi - self.lo) + self.lo

    def wide(self, v):
        return v.reshape((-1,) + (1,) * (self.x.ndim - 1))

    def runner_up(self, z, onehot):
        return z * (1 - onehot) + (np.min(z, axis=1) - 1)[:, None] * onehot

    def loss(self, rows, adv, c):
        """(logits, squared L2 distance, c * hinge + distance) for the rows `rows` (a boolean mask) moved to `adv`."""
        l2 = np.sum(np.square(self.x[rows] - adv).reshape(adv.shape[0], -1), axis=1)
        z = self.eng.logits(adv.astype(np.float32)).astype(np.float64)
        onehot = self.onehot[rows]
        hinge = np.maximum(np.sum(z * onehot, axis=1) - np.max(self.runner_up(z, onehot), axis=1) + self.cfg["confidence"], 0.0)
        return z, l2, c[rows] * hinge + l2

    def descent_direction(self, rows, z, adv, adv_tanh, c):
        onehot = self.onehot[rows]
        head = onehot.copy()
        head[np.arange(head.shape[0]), np.argmax(self.runner_up(z[rows], onehot), axis=1)] -= 1.0
        grad, _ = self.eng.logit_gradient(adv[rows].astype(np.float32), head.astype(np.float32))
        g = grad.cpu().numpy().astype(np.float64) * self.wide(c[rows]) + 2 * (adv[rows] - self.x[rows])
        return -g * (self.hi - self.lo) * (1 - np.square(np.tanh(adv_tanh[rows]))) / (2 * _SMOOTH)

    def keep_best(self, success, l2, adv):
        better = success & (l2 < self.best_l2)
        self.best_l2[better] = l2[better]
        self.best_adv[better] = adv[better]

    def search_step(self, c):
        """One binary-search step at constants c: max_iter line-searched descent steps.  Returns the rows that were
        adversarial at any point."""
        cfg, n = self.cfg, self.n
        everyone = np.ones(n, bool)
        lr = cfg["learning_rate"] * np.ones(n)
        adv, adv_tanh = self.x.copy(), self.x_tanh.copy()
        z, l2, loss = self.loss(everyone, adv, c)
        success = loss - l2 <= 0
        ever = success
        for _ in range(cfg["max_iter"]):
            self.keep_best(success, l2, adv)
            active = (c < _C_LIMIT) & (lr > 0)
            if not active.any():
                break
            step = self.descent_direction(active, z, adv, adv_tanh, c)
            before, lowest = loss.copy(), loss.copy()
            lowest_lr, halved = np.zeros(n), np.zeros(n)

            def probe(which):                         # `which`: mask over the ACTIVE rows
                rows = active.copy()
                rows[active] = which
                trial = self.image(adv_tanh[rows] + self.wide(lr[rows]) * step[which])
                _, l2[rows], loss[rows] = self.loss(rows, trial, c)
                lower = loss < lowest
                lowest_lr[lower] = lr[lower]
                lowest[lower] = loss[lower]
                return rows

            for _ in range(cfg["max_halving"]):
                which = loss[active] >= before[active]
                if not which.any():
                    break
                rows = probe(which)
                lr[rows] /= 2
                halved[rows] += 1
            lr[active] *= 2
            for _ in range(cfg["max_doubling"]):
                which = (halved[active] == 1) & (loss[active] <= lowest[active])
                if not which.any():
                    break
                rows = active.copy()
                rows[active] = which
                lr[rows] *= 2
                probe(which)
            lr[halved == 1] /= 2
            which = lowest_lr[active] > 0
            if which.any():
                rows = active.copy()
                rows[active] = which
                adv_tanh[rows] = adv_tanh[rows] + self.wide(lowest_lr[rows]) * step[which]
                adv[rows] = self.image(adv_tanh[rows])
                z[rows], l2[rows], loss[rows] = self.loss(rows, adv[rows], c)
                success = loss - l2 <= 0
                ever = ever | success
        self.keep_best(success, l2, adv)
        return ever

    def run(self):
        cfg, n = self.cfg, self.n
        c = cfg["initial_const"] * np.ones(n)
        floor = np.zeros(n)
        doubling = np.ones(n, bool)
        for _ in range(cfg["binary_search_steps"]):
            if not (c < _C_LIMIT).any():
                break
            ever = self.search_step(c)
            doubling[ever] = False
            c[ever] = (floor + c)[ever] / 2
            c[~ever & doubling] *= 2
            c[~ever & ~doubling] += (c - floor)[~ever & ~doubling] / 2
            floor[~ever] = c[~ever]                 # ART keeps an alias of c here, so the floor follows the NEW constant
        return self.best_adv.astype(np.float32)


def carlini_l2(net, x, params=None):
    """Untargeted Carlini-Wagner L2 examples for the host array ``x``; same shape and dtype float32."""
    cfg = dict(CW_DEFAULTS)
    cfg.update({k: v for k, v in (params or {}).items() if k in cfg})
    x = np.ascontiguousarray(x, dtype=np.float32)
    adv = np.empty_like(x)
    for lo in range(0, x.shape[0], int(cfg["batch_size"])):
        xb = x[lo: lo + int(cfg["batch_size"])]
        adv[lo: lo + xb.shape[0]] = _CarliniBatch(net, xb, cfg).run()
    return adv
