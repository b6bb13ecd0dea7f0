"""Robustness metrics under the reference's names (/root/reference/model/metrics.py:57-91, 205-324).

The reference's module is IBM ART 0.9.0 code driven by a KerasClassifier; here the same names resolve to the
implementations that run on the CUDA engine (featurenet_b200/robustness.py: gradient attacks, CLEVER, Carlini-Wagner).
The first argument is the trained network (what ``TensorflowGenerator(...).model.model`` holds) instead of an ART
classifier wrapper; the remaining arguments keep the reference's order and meaning.

One deliberate difference, also recorded in DESIGN.md / INTEGRATION.md: when no row changes its prediction the
reference's ``empirical_robustness`` returns a bare ``0`` (model/metrics.py:79), which makes the tuple-unpacking at
its only call site raise inside ``eval_robustness``'s try block and skips the remaining scores of that candidate.
Here ``(0.0, x_adv)`` is returned and scoring continues.
"""
from featurenet_b200 import robustness as _r


def empirical_robustness(classifier, x, attack_name, attack_params=None):
    return _r.empirical_robustness(classifier, x, attack_name, attack_params)


def clever_u(classifier, x, nb_batches, batch_size, radius, norm, c_init=1, pool_factor=10):
    return _r.clever_u(classifier, x, nb_batches, batch_size, radius, norm=norm, c_init=c_init, pool_factor=pool_factor)


def clever_t(classifier, x, target_class, nb_batches, batch_size, radius, norm, c_init=1, pool_factor=10):
    """Targeted CLEVER score of one sample for one target class (model/metrics.py:242-324)."""
    return _r.clever_t(classifier, x, target_class, nb_batches, batch_size, radius, norm=norm, c_init=c_init,
                       pool_factor=pool_factor)


def clever(classifier, x, nb_batches, batch_size, radius, norm, target=None, target_sort=False, c_init=1, pool_factor=10):
    """model/metrics.py:153-202: per-class targeted scores (None at the predicted class)."""
    import numpy as np
    pred = int(np.argmax(classifier.logits(np.asarray([x], dtype=np.float32)), axis=1)[0])
    classes = classifier.output_shape[-1]
    if target is None:
        targets = np.argsort(classifier.logits(np.asarray([x], dtype=np.float32)))[0][::-1] if target_sort else range(classes)
    elif isinstance(target, (int, np.integer)):
        targets = [target]
    else:
        targets = target
    return np.array([None if j == pred else clever_t(classifier, x, j, nb_batches, batch_size, radius, norm, c_init, pool_factor)
                     for j in targets])
