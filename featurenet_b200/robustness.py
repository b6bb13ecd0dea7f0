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
run in one kernel per step (``fnet_l2_attack_step``).  ``cw`` (Carlini-Wagner L2) and ``clever`` are not built.
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
    if attack not in ("fgsm", "pgd"):
        raise NotImplementedError("attack %r is not built (fgsm and pgd are)" % attack)
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
