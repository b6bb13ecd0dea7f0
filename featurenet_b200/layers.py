"""Symbolic layer constructors: the Keras surface that FeatureNet's graph description calls.

The reference builds its candidates by calling ``keras.layers.*`` on tensors
(/root/reference/model/input.py:190,233,237,240,294-298, model/operation.py:93,114,136,150,
165,186,222,238,248, model/keras_model.py:108,122-126).  Here the same calls record a layer
DAG (``LayerNode``) with inferred NHWC shapes and weight specs; ``Model(...)`` prunes it to
the layers reachable from the output, exactly what ``keras.models.Model`` keeps, and hands
it to the CUDA executor in ``featurenet_b200.network``.

Shape inference restates Keras 2.2.4 / TF 1.12 rules (third-party, restated from the
published behaviour): SAME => out = ceil(in / stride); VALID => out = floor((in - k_eff) /
stride) + 1 and a non-positive size is an error, as in TF.
"""
import math

ACTIVATIONS = (None, "linear", "relu", "tanh", "sigmoid", "softmax")

_uid_counters = {}


def _auto_name(prefix):
    n = _uid_counters.get(prefix, 0) + 1
    _uid_counters[prefix] = n
    return "%s_%d" % (prefix, n)


def reset_uids():
    """Counterpart of keras.backend.clear_session() for the auto-naming counters."""
    _uid_counters.clear()


class Dimension(object):
    """One entry of a static shape; mirrors tf.Dimension far enough for the graph code."""
    __slots__ = ("value",)

    def __init__(self, value):
        self.value = value

    def _v(self, other):
        return other.value if isinstance(other, Dimension) else other

    def __eq__(self, other):
        return self.value == self._v(other)

    def __ne__(self, other):
        return self.value != self._v(other)

    def __lt__(self, other):
        return self.value < self._v(other)

    def __le__(self, other):
        return self.value <= self._v(other)

    def __gt__(self, other):
        return self.value > self._v(other)

    def __ge__(self, other):
        return self.value >= self._v(other)

    def __hash__(self):
        return hash(self.value)

    def __int__(self):
        return int(self.value)

    def __repr__(self):
        return "Dimension(%r)" % (self.value,)


class TensorShape(object):
    """Static shape with an unknown (None) batch axis."""

    def __init__(self, dims):
        self._dims = tuple(None if d is None else int(d) for d in dims)

    @property
    def ndims(self):
        return len(self._dims)

    @property
    def dims(self):
        return [Dimension(d) for d in self._dims]

    def as_list(self):
        return list(self._dims)

    def __len__(self):
        return len(self._dims)

    def __iter__(self):
        return iter(self._dims)

    def __getitem__(self, i):
        if isinstance(i, slice):
            return TensorShape(self._dims[i])
        return Dimension(self._dims[i])

    def __eq__(self, other):
        return self.as_list() == (other.as_list() if isinstance(other, TensorShape) else list(other))

    def __repr__(self):
        return "TensorShape(%r)" % (list(self._dims),)


class SymTensor(object):
    """Output of a recorded layer."""

    def __init__(self, shape, node):
        self.shape = TensorShape(shape)
        self.node = node
        self.name = node.name + "/out"

    @property
    def static(self):
        """Shape without the batch axis, as a tuple of ints."""
        return tuple(self.shape.as_list()[1:])

    def __repr__(self):
        return "<SymTensor %s %r>" % (self.name, self.shape.as_list())


class WeightSpec(object):
    __slots__ = ("name", "shape", "init", "trainable", "fans")

    def __init__(self, name, shape, init, trainable=True, fans=None):
        self.name = name
        self.shape = tuple(int(s) for s in shape)
        self.init = init            # 'glorot_uniform' | 'zeros' | 'ones'
        self.trainable = trainable
        self.fans = fans            # (fan_in, fan_out) for glorot

    @property
    def size(self):
        n = 1
        for s in self.shape:
            n *= s
        return n


class LayerNode(object):
    """One recorded layer application (layers are never shared in FeatureNet graphs)."""

    def __init__(self, kind, name, cfg, inputs, out_shape, weights=()):
        self.kind = kind
        self.name = name
        self.cfg = dict(cfg)
        self.inputs = list(inputs)          # SymTensor list
        self.weights = list(weights)
        self.output = SymTensor(out_shape, self)

    @property
    def nb_params(self):
        return sum(w.size for w in self.weights)

    @property
    def class_name(self):
        return _KIND_TO_KERAS[self.kind]

    def describe(self):
        """JSON-able summary used by the structural parity fixtures."""
        return {
            "class": self.class_name,
            "cfg": {k: (list(v) if isinstance(v, tuple) else v) for k, v in sorted(self.cfg.items())},
            "in_shapes": [t.shape.as_list() for t in self.inputs],
            "out_shape": self.output.shape.as_list(),
            "params": self.nb_params,
        }

    def __repr__(self):
        return "<%s %s -> %r>" % (self.class_name, self.name, self.output.shape.as_list())


_KIND_TO_KERAS = {
    "input": "InputLayer", "conv2d": "Conv2D", "sepconv2d": "SeparableConv2D",
    "dwconv2d": "DepthwiseConv2D", "dense": "Dense", "maxpool": "MaxPooling2D",
    "avgpool": "AveragePooling2D", "gap": "GlobalAveragePooling2D", "flatten": "Flatten",
    "dropout": "Dropout", "bn": "BatchNormalization", "act": "Activation", "add": "Add",
    "mul": "Multiply", "concat": "Concatenate", "pad": "ZeroPadding2D",
}


# ----------------------------------------------------------------------------- helpers
def _pair(v):
    if isinstance(v, (tuple, list)):
        if len(v) != 2:
            raise ValueError("expected a pair, got %r" % (v,))
        return (int(v[0]), int(v[1]))
    return (int(v), int(v))


def _check_act(a):
    if a not in ACTIVATIONS:
        raise ValueError("unknown activation %r" % (a,))
    return None if a == "linear" else a


def _check_padding(p):
    p = str(p).lower()
    if p not in ("same", "valid"):
        raise ValueError("unknown padding %r" % (p,))
    return p


def conv_out_len(n, k, s, padding, d=1):
    """Output length and leading pad along one axis (TF rule; extra pad goes after)."""
    k_eff = (k - 1) * d + 1
    if padding == "same":
        out = -(-n // s)
        total = max((out - 1) * s + k_eff - n, 0)
        return out, total // 2
    out = (n - k_eff) // s + 1 if n >= k_eff else 0
    if out <= 0:
        raise ValueError("negative dimension: input %d, kernel %d, padding valid" % (n, k_eff))
    return out, 0


def _need_rank(x, rank, who):
    if x.shape.ndims != rank:
        raise ValueError("%s expects a rank-%d input, got %r" % (who, rank, x.shape.as_list()))


class _Layer(object):
    prefix = "layer"

    def __init__(self, name=None):
        self.name = name if name else _auto_name(self.prefix)

    def __call__(self, x):
        raise NotImplementedError


# ----------------------------------------------------------------------------- layers
def Input(shape=None, name=None, **_):
    """keras.layers.Input: the placeholder the candidate consumes (keras_model.py:108)."""
    name = name if name else _auto_name("input")
    node = LayerNode("input", name, {}, [], (None,) + tuple(int(s) for s in shape))
    return node.output


class Conv2D(_Layer):
    """NHWC convolution, HWIO kernel, bias, fused activation (input.py:294, operation.py:186)."""
    prefix = "conv2d"

    def __init__(self, filters, kernel_size, strides=(1, 1), padding="valid", activation=None,
                 dilation_rate=(1, 1), name=None, use_bias=True, **_):
        super(Conv2D, self).__init__(name)
        self.filters = int(filters)
        self.kernel = _pair(kernel_size)
        self.strides = _pair(strides)
        self.dilation = _pair(dilation_rate)
        self.padding = _check_padding(padding)
        self.activation = _check_act(activation)
        self.use_bias = use_bias
        if self.filters <= 0 or min(self.kernel) <= 0 or min(self.strides) <= 0:
            raise ValueError("bad Conv2D configuration")
        if max(self.strides) > 1 and max(self.dilation) > 1:
            raise ValueError("strides > 1 not supported in conjunction with dilation_rate > 1")

    def __call__(self, x):
        _need_rank(x, 4, "Conv2D")
        _, h, w, c = x.shape.as_list()
        ho, pt = conv_out_len(h, self.kernel[0], self.strides[0], self.padding, self.dilation[0])
        wo, pl = conv_out_len(w, self.kernel[1], self.strides[1], self.padding, self.dilation[1])
        kh, kw = self.kernel
        ws = [WeightSpec("kernel", (kh, kw, c, self.filters), "glorot_uniform",
                         fans=(kh * kw * c, kh * kw * self.filters))]
        if self.use_bias:
            ws.append(WeightSpec("bias", (self.filters,), "zeros"))
        cfg = dict(filters=self.filters, kernel=self.kernel, strides=self.strides,
                   padding=self.padding, activation=self.activation, dilation=self.dilation,
                   pad_top=pt, pad_left=pl)
        return LayerNode("conv2d", self.name, cfg, [x], (None, ho, wo, self.filters), ws).output


class DepthwiseConv2D(_Layer):
    """Depthwise convolution, depth multiplier 1, bias, activation (input.py:298)."""
    prefix = "depthwise_conv2d"

    def __init__(self, kernel_size, strides=(1, 1), padding="valid", activation=None, name=None, **_):
        super(DepthwiseConv2D, self).__init__(name)
        self.kernel = _pair(kernel_size)
        self.strides = _pair(strides)
        self.padding = _check_padding(padding)
        self.activation = _check_act(activation)

    def __call__(self, x):
        _need_rank(x, 4, "DepthwiseConv2D")
        _, h, w, c = x.shape.as_list()
        kh, kw = self.kernel
        ho, pt = conv_out_len(h, kh, self.strides[0], self.padding)
        wo, pl = conv_out_len(w, kw, self.strides[1], self.padding)
        ws = [WeightSpec("depthwise_kernel", (kh, kw, c, 1), "glorot_uniform", fans=(kh * kw * c, kh * kw)),
              WeightSpec("bias", (c,), "zeros")]
        cfg = dict(kernel=self.kernel, strides=self.strides, padding=self.padding,
                   activation=self.activation, pad_top=pt, pad_left=pl)
        return LayerNode("dwconv2d", self.name, cfg, [x], (None, ho, wo, c), ws).output


class SeparableConv2D(_Layer):
    """Depthwise (no bias, no activation) then pointwise 1x1 + bias + activation (input.py:296)."""
    prefix = "separable_conv2d"

    def __init__(self, filters, kernel_size, strides=(1, 1), padding="valid", activation=None, name=None, **_):
        super(SeparableConv2D, self).__init__(name)
        self.filters = int(filters)
        self.kernel = _pair(kernel_size)
        self.strides = _pair(strides)
        self.padding = _check_padding(padding)
        self.activation = _check_act(activation)

    def __call__(self, x):
        _need_rank(x, 4, "SeparableConv2D")
        _, h, w, c = x.shape.as_list()
        kh, kw = self.kernel
        ho, pt = conv_out_len(h, kh, self.strides[0], self.padding)
        wo, pl = conv_out_len(w, kw, self.strides[1], self.padding)
        ws = [WeightSpec("depthwise_kernel", (kh, kw, c, 1), "glorot_uniform", fans=(kh * kw * c, kh * kw)),
              WeightSpec("pointwise_kernel", (1, 1, c, self.filters), "glorot_uniform", fans=(c, self.filters)),
              WeightSpec("bias", (self.filters,), "zeros")]
        cfg = dict(filters=self.filters, kernel=self.kernel, strides=self.strides, padding=self.padding,
                   activation=self.activation, pad_top=pt, pad_left=pl)
        return LayerNode("sepconv2d", self.name, cfg, [x], (None, ho, wo, self.filters), ws).output


def Conv1D(*a, **k):
    raise NotImplementedError("Conv1D: rank-3 tensors never arise from NHWC image inputs (input.py:301-306)")


SeparableConv1D = Conv1D


class Dense(_Layer):
    """act(x . W + b) over the last axis of a rank-2 or rank-4 tensor (input.py:190, keras_model.py:124)."""
    prefix = "dense"

    def __init__(self, units, activation=None, name=None, **_):
        super(Dense, self).__init__(name)
        self.units = int(units)
        self.activation = _check_act(activation)
        if self.units <= 0:
            raise ValueError("Dense units must be positive")

    def __call__(self, x):
        if x.shape.ndims < 2:
            raise ValueError("Dense needs rank >= 2")
        shp = x.shape.as_list()
        k = shp[-1]
        ws = [WeightSpec("kernel", (k, self.units), "glorot_uniform", fans=(k, self.units)),
              WeightSpec("bias", (self.units,), "zeros")]
        cfg = dict(units=self.units, activation=self.activation)
        return LayerNode("dense", self.name, cfg, [x], tuple(shp[:-1]) + (self.units,), ws).output


class _Pool2D(_Layer):
    kind = None

    def __init__(self, pool_size=(2, 2), strides=None, padding="valid", name=None, **_):
        super(_Pool2D, self).__init__(name)
        self.pool = _pair(pool_size)
        self.strides = _pair(strides) if strides is not None else self.pool
        self.padding = _check_padding(padding)
        if min(self.pool) <= 0 or min(self.strides) <= 0:
            raise ValueError("bad pooling configuration")

    def __call__(self, x):
        _need_rank(x, 4, self.__class__.__name__)
        _, h, w, c = x.shape.as_list()
        ho, pt = conv_out_len(h, self.pool[0], self.strides[0], self.padding)
        wo, pl = conv_out_len(w, self.pool[1], self.strides[1], self.padding)
        cfg = dict(pool=self.pool, strides=self.strides, padding=self.padding, pad_top=pt, pad_left=pl)
        return LayerNode(self.kind, self.name, cfg, [x], (None, ho, wo, c)).output


class MaxPooling2D(_Pool2D):
    prefix = "max_pooling2d"
    kind = "maxpool"


class AveragePooling2D(_Pool2D):
    prefix = "average_pooling2d"
    kind = "avgpool"


class GlobalAveragePooling2D(_Layer):
    prefix = "global_average_pooling2d"

    def __call__(self, x):
        _need_rank(x, 4, "GlobalAveragePooling2D")
        return LayerNode("gap", self.name, {}, [x], (None, x.shape.as_list()[3])).output


class Flatten(_Layer):
    prefix = "flatten"

    def __call__(self, x):
        n = 1
        for d in x.shape.as_list()[1:]:
            n *= d
        return LayerNode("flatten", self.name, {}, [x], (None, n)).output


class Dropout(_Layer):
    prefix = "dropout"

    def __init__(self, rate, name=None, **_):
        super(Dropout, self).__init__(name)
        self.rate = float(rate)
        if not 0.0 <= self.rate < 1.0:
            raise ValueError("dropout rate must be in [0, 1)")

    def __call__(self, x):
        return LayerNode("dropout", self.name, dict(rate=self.rate), [x], x.shape.as_list()).output


class BatchNormalization(_Layer):
    """Keras BN; FeatureNet hard-wires axis=1 (operation.py:142), i.e. the H axis of NHWC."""
    prefix = "batch_normalization"

    def __init__(self, axis=-1, momentum=0.99, epsilon=1e-3, name=None, **_):
        super(BatchNormalization, self).__init__(name)
        self.axis = int(axis)
        self.momentum = float(momentum)
        self.epsilon = float(epsilon)

    def __call__(self, x):
        shp = x.shape.as_list()
        axis = self.axis if self.axis >= 0 else len(shp) + self.axis
        if axis <= 0 or axis >= len(shp):
            raise ValueError("bad BatchNormalization axis %d for rank %d" % (self.axis, len(shp)))
        n = shp[axis]
        ws = [WeightSpec("gamma", (n,), "ones"), WeightSpec("beta", (n,), "zeros"),
              WeightSpec("moving_mean", (n,), "zeros", trainable=False),
              WeightSpec("moving_variance", (n,), "ones", trainable=False)]
        cfg = dict(axis=axis, momentum=self.momentum, epsilon=self.epsilon)
        return LayerNode("bn", self.name, cfg, [x], shp, ws).output


class Activation(_Layer):
    prefix = "activation"

    def __init__(self, activation, name=None, **_):
        super(Activation, self).__init__(name)
        self.activation = _check_act(activation)

    def __call__(self, x):
        return LayerNode("act", self.name, dict(activation=self.activation), [x], x.shape.as_list()).output


class ZeroPadding2D(_Layer):
    prefix = "zero_padding2d"

    def __init__(self, padding=(1, 1), name=None, **_):
        super(ZeroPadding2D, self).__init__(name)
        self.pad = _pair(padding)
        if min(self.pad) < 0:
            raise ValueError("negative padding")

    def __call__(self, x):
        _need_rank(x, 4, "ZeroPadding2D")
        _, h, w, c = x.shape.as_list()
        cfg = dict(pad=self.pad)
        return LayerNode("pad", self.name, cfg, [x], (None, h + 2 * self.pad[0], w + 2 * self.pad[1], c)).output


class _Merge(_Layer):
    kind = None

    def __call__(self, xs):
        xs = list(xs)
        if len(xs) != 2:
            raise ValueError("merge layers take exactly two inputs here")
        a, b = xs[0].shape.as_list(), xs[1].shape.as_list()
        if a != b:
            # Keras would try to broadcast size-1 axes; FeatureNet graphs never rely on that
            raise ValueError("Operands could not be broadcast together with shapes %r %r" % (a, b))
        return LayerNode(self.kind, self.name, {}, xs, a).output


class Add(_Merge):
    prefix = "add"
    kind = "add"


class Multiply(_Merge):
    prefix = "multiply"
    kind = "mul"


class Concatenate(_Layer):
    """Concatenate; FeatureNet hard-wires axis=1 (operation.py:228): stacks along H."""
    prefix = "concatenate"

    def __init__(self, axis=-1, name=None, **_):
        super(Concatenate, self).__init__(name)
        self.axis = int(axis)

    def __call__(self, xs):
        xs = list(xs)
        if len(xs) != 2:
            raise ValueError("Concatenate takes exactly two inputs here")
        a, b = xs[0].shape.as_list(), xs[1].shape.as_list()
        axis = self.axis if self.axis >= 0 else len(a) + self.axis
        if len(a) != len(b) or axis <= 0 or axis >= len(a):
            raise ValueError("bad Concatenate inputs %r %r axis %d" % (a, b, self.axis))
        for i in range(1, len(a)):
            if i != axis and a[i] != b[i]:
                raise ValueError("Concatenate shapes differ off-axis: %r %r" % (a, b))
        out = list(a)
        out[axis] = a[axis] + b[axis]
        return LayerNode("concat", self.name, dict(axis=axis), xs, out).output


# ----------------------------------------------------------------------------- graph
def reachable_layers(outputs):
    """Layers that feed ``outputs`` in a deterministic topological order (inputs first)."""
    order, seen = [], set()

    def visit(node):
        if id(node) in seen:
            return
        seen.add(id(node))
        for t in node.inputs:
            visit(t.node)
        order.append(node)

    # iterative to survive very deep mutated candidates
    import sys
    limit = sys.getrecursionlimit()
    sys.setrecursionlimit(max(limit, 20000))
    try:
        for t in outputs:
            visit(t.node)
    finally:
        sys.setrecursionlimit(limit)
    return order


def Model(inputs=None, outputs=None, name=None):
    """keras.models.Model: keep the layers reachable from the outputs (keras_model.py:126)."""
    from .network import Network
    outs = outputs if isinstance(outputs, (list, tuple)) else [outputs]
    return Network(inputs, outs[0], name=name)


def glorot_limit(fans):
    return math.sqrt(6.0 / float(fans[0] + fans[1]))
