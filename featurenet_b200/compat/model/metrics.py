"""Robustness metrics (CLEVER, empirical robustness) are SURVEY.md section 8(f) item 1: not built yet.

The reference's model/metrics.py is a copy of IBM ART 0.9.0 code driven by FGSM/PGD/C&W attacks;
it needs input gradients from the engine and the ``art`` package, neither of which is in scope
for the per-candidate train+evaluate path.
"""


def _unavailable(*_, **__):
    raise NotImplementedError("robustness metrics are outside the train+evaluate hot path (SURVEY.md 8f.1)")


empirical_robustness = clever_u = clever_t = loss_sensitivity = _unavailable
