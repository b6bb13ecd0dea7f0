"""Optimizer descriptions (hyper-parameters only; the update itself is one CUDA launch, csrc/optim.cu).

Mirrors the two optimizers the reference can name at tensorflow_generator.py:232:
``Adam(lr=lr_schedule(0))`` (the one every candidate is compiled with) and the string "sgd".
"""


class Optimizer(object):
    kind = None

    def __init__(self, lr):
        self.lr = float(lr)


class Adam(Optimizer):
    """keras.optimizers.Adam defaults of Keras 2.2.4: beta_1 0.9, beta_2 0.999, epsilon K.epsilon()=1e-7, no decay."""
    kind = "adam"

    def __init__(self, lr=0.001, beta_1=0.9, beta_2=0.999, epsilon=None, **_):
        super(Adam, self).__init__(lr)
        self.beta_1, self.beta_2 = float(beta_1), float(beta_2)
        self.epsilon = 1e-7 if epsilon is None else float(epsilon)


class SGD(Optimizer):
    """keras.optimizers.SGD: lr 0.01, momentum 0, no nesterov."""
    kind = "sgd"

    def __init__(self, lr=0.01, momentum=0.0, nesterov=False, **_):
        super(SGD, self).__init__(lr)
        self.momentum, self.nesterov = float(momentum), bool(nesterov)


def resolve_optimizer(spec):
    if spec is None:
        return Adam()
    if isinstance(spec, Optimizer):
        return spec
    if isinstance(spec, str):
        name = spec.lower()
        if name == "adam":
            return Adam()
        if name == "sgd":
            return SGD()
    raise ValueError("unknown optimizer %r" % (spec,))
