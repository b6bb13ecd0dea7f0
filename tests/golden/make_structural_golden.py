#!/usr/bin/env python
"""Generate tests/golden/structural.json by executing the REFERENCE's own graph code.

FeatureNet's arithmetic lives in Keras/TensorFlow (not installed, not installable here), but
its graph construction -- /root/reference/model/*.py, model/mutation/*.py, Products_tree.py --
is plain Python.  This script imports those files read-only from /root/reference with tiny
shape-only stand-ins for ``keras`` / ``tensorflow`` / ``validators`` registered in
``sys.modules``, builds a set of candidates (templates, sampled products, the shipped
10Products.pdt, seeded mutation walks, breeding) and records for each the layer list the
reference emits: class, configuration, shapes, parameter count and connectivity.

The stand-ins here are deliberately independent of featurenet_b200.layers (the product's
symbolic layers): tests/test_structural_parity.py compares the two.

Run in the build container only (needs /root/reference):
    python tests/golden/make_structural_golden.py
"""
import copy
import json
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
REFERENCE = os.environ.get("FNET_REFERENCE", "/root/reference")


# ----------------------------------------------------------------------------- keras stand-ins
class Dim(object):
    def __init__(self, v):
        self.value = v

    def __lt__(self, o):
        return self.value < o.value

    def __eq__(self, o):
        return self.value == (o.value if isinstance(o, Dim) else o)

    def __ne__(self, o):
        return not self.__eq__(o)

    def __hash__(self):
        return hash(self.value)


class Shape(object):
    def __init__(self, dims):
        self._d = list(dims)
        self.ndims = len(self._d)
        self.dims = [Dim(v) for v in self._d]

    def as_list(self):
        return list(self._d)


class T(object):
    count = 0

    def __init__(self, shape, layer, inbound):
        T.count += 1
        self.shape = Shape(shape)
        self.layer = layer
        self.inbound = inbound
        self.name = "t%d" % T.count


def _pair(v):
    if isinstance(v, (tuple, list)):
        return [int(v[0]), int(v[1])]
    return [int(v), int(v)]


def _out_len(n, k, s, padding):
    if padding == "same":
        return (n + s - 1) // s
    if padding == "valid":
        o = (n - k) // s + 1
        if n < k or o <= 0:
            raise ValueError("Negative dimension size")
        return o
    raise ValueError("bad padding %r" % (padding,))


class Layer(object):
    created = 0

    def __init__(self, **kw):
        Layer.created += 1
        self.name = kw.get("name")
        self.params = 0
        self.cfg = {}

    def __call__(self, x):
        xs = x if isinstance(x, list) else [x]
        out_shape = self.compute(xs)
        return T(out_shape, self, xs)


def _act(a):
    if a not in (None, "relu", "tanh", "sigmoid", "softmax", "linear"):
        raise ValueError("unknown activation %r" % (a,))
    return a


class Conv2D(Layer):
    def __init__(self, filters, kernel_size, strides=(1, 1), padding="valid", activation=None, dilation_rate=(1, 1),
                 name=None):
        Layer.__init__(self, name=name)
        self.cfg = dict(filters=int(filters), kernel=_pair(kernel_size), strides=_pair(strides), padding=padding,
                        activation=_act(activation), dilation=_pair(dilation_rate))

    def compute(self, xs):
        (x,) = xs
        b, h, w, c = x.shape.as_list()
        kh, kw = self.cfg["kernel"]
        dh, dw = self.cfg["dilation"]
        sh, sw = self.cfg["strides"]
        self.params = kh * kw * c * self.cfg["filters"] + self.cfg["filters"]
        return [b, _out_len(h, (kh - 1) * dh + 1, sh, self.cfg["padding"]),
                _out_len(w, (kw - 1) * dw + 1, sw, self.cfg["padding"]), self.cfg["filters"]]


class SeparableConv2D(Layer):
    def __init__(self, filters, kernel_size, strides=(1, 1), padding="valid", activation=None, name=None):
        Layer.__init__(self, name=name)
        self.cfg = dict(filters=int(filters), kernel=_pair(kernel_size), strides=_pair(strides), padding=padding,
                        activation=_act(activation))

    def compute(self, xs):
        (x,) = xs
        b, h, w, c = x.shape.as_list()
        kh, kw = self.cfg["kernel"]
        sh, sw = self.cfg["strides"]
        f = self.cfg["filters"]
        self.params = kh * kw * c + c * f + f
        return [b, _out_len(h, kh, sh, self.cfg["padding"]), _out_len(w, kw, sw, self.cfg["padding"]), f]


class DepthwiseConv2D(Layer):
    def __init__(self, kernel_size, strides=(1, 1), padding="valid", activation=None, name=None):
        Layer.__init__(self, name=name)
        self.cfg = dict(kernel=_pair(kernel_size), strides=_pair(strides), padding=padding, activation=_act(activation))

    def compute(self, xs):
        (x,) = xs
        b, h, w, c = x.shape.as_list()
        kh, kw = self.cfg["kernel"]
        sh, sw = self.cfg["strides"]
        self.params = kh * kw * c + c
        return [b, _out_len(h, kh, sh, self.cfg["padding"]), _out_len(w, kw, sw, self.cfg["padding"]), c]


class _NoRank3(Layer):
    def __init__(self, *a, **k):
        raise NotImplementedError("rank-3 convolution")


Conv1D = SeparableConv1D = _NoRank3


class Dense(Layer):
    def __init__(self, units, activation=None, name=None):
        Layer.__init__(self, name=name)
        self.cfg = dict(units=int(units), activation=_act(activation))

    def compute(self, xs):
        (x,) = xs
        shp = x.shape.as_list()
        self.params = shp[-1] * self.cfg["units"] + self.cfg["units"]
        return shp[:-1] + [self.cfg["units"]]


class _Pool(Layer):
    def __init__(self, pool_size=(2, 2), strides=None, padding="valid", name=None):
        Layer.__init__(self, name=name)
        pool = _pair(pool_size)
        self.cfg = dict(pool=pool, strides=_pair(strides) if strides is not None else pool, padding=padding)

    def compute(self, xs):
        (x,) = xs
        b, h, w, c = x.shape.as_list()
        return [b, _out_len(h, self.cfg["pool"][0], self.cfg["strides"][0], self.cfg["padding"]),
                _out_len(w, self.cfg["pool"][1], self.cfg["strides"][1], self.cfg["padding"]), c]


class MaxPooling2D(_Pool):
    pass


class AveragePooling2D(_Pool):
    pass


class GlobalAveragePooling2D(Layer):
    def __init__(self, name=None):
        Layer.__init__(self, name=name)

    def compute(self, xs):
        b, h, w, c = xs[0].shape.as_list()
        return [b, c]


class Flatten(Layer):
    def __init__(self, name=None):
        Layer.__init__(self, name=name)

    def compute(self, xs):
        n = 1
        for d in xs[0].shape.as_list()[1:]:
            n *= d
        return [None, n]


class Dropout(Layer):
    def __init__(self, rate, name=None):
        Layer.__init__(self, name=name)
        self.cfg = dict(rate=float(rate))
        if not 0 <= float(rate) < 1:
            raise ValueError("rate")

    def compute(self, xs):
        return xs[0].shape.as_list()


class BatchNormalization(Layer):
    def __init__(self, axis=-1, name=None):
        Layer.__init__(self, name=name)
        self.cfg = dict(axis=int(axis))

    def compute(self, xs):
        shp = xs[0].shape.as_list()
        self.params = 4 * shp[self.cfg["axis"]]
        return shp


class Activation(Layer):
    def __init__(self, activation=None, name=None):
        Layer.__init__(self, name=name)
        self.cfg = dict(activation=_act(activation))

    def compute(self, xs):
        return xs[0].shape.as_list()


class ZeroPadding2D(Layer):
    def __init__(self, padding=(1, 1), name=None):
        Layer.__init__(self, name=name)
        self.cfg = dict(pad=_pair(padding))

    def compute(self, xs):
        b, h, w, c = xs[0].shape.as_list()
        return [b, h + 2 * self.cfg["pad"][0], w + 2 * self.cfg["pad"][1], c]


class _Same(Layer):
    def __init__(self, name=None):
        Layer.__init__(self, name=name)

    def compute(self, xs):
        a, b = xs[0].shape.as_list(), xs[1].shape.as_list()
        if a != b:
            raise ValueError("Operands could not be broadcast together")
        return a


class Add(_Same):
    pass


class Multiply(_Same):
    pass


class Concatenate(Layer):
    def __init__(self, axis=-1, name=None):
        Layer.__init__(self, name=name)
        self.cfg = dict(axis=int(axis))

    def compute(self, xs):
        a, b = xs[0].shape.as_list(), xs[1].shape.as_list()
        ax = self.cfg["axis"]
        for i in range(1, len(a)):
            if i != ax and a[i] != b[i]:
                raise ValueError("concat shapes")
        out = list(a)
        out[ax] = a[ax] + b[ax]
        return out


class InputLayer(Layer):
    def __init__(self):
        Layer.__init__(self)


def Input(shape):
    return T([None] + list(shape), InputLayer(), [])


class Model(object):
    def __init__(self, outputs=None, inputs=None, name=None):
        outs = outputs if isinstance(outputs, list) else [outputs]
        order, seen = [], set()

        def visit(t):
            if id(t) in seen:
                return
            seen.add(id(t))
            for src in t.inbound:
                visit(src)
            order.append(t)

        sys.setrecursionlimit(20000)
        visit(outs[0])
        self.tensors = order
        self.layers = [t.layer for t in order]
        self.name = name

    def count_params(self):
        return sum(l.params for l in self.layers)

    def summary(self):
        pass

    def describe(self):
        pos = {id(t): i for i, t in enumerate(self.tensors)}
        rows = []
        for t in self.tensors:
            rows.append({"class": type(t.layer).__name__, "cfg": t.layer.cfg, "out_shape": t.shape.as_list(),
                         "params": t.layer.params, "inbound": [pos[id(s)] for s in t.inbound]})
        return rows


def _install_stubs():
    def mod(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    def zeros(shape):
        raise ValueError("Cannot convert a partially known TensorShape to a Tensor")

    backend = mod("keras.backend", zeros=zeros, image_data_format=lambda: "channels_last")
    layers = mod("keras.layers", **{c.__name__: c for c in (
        Conv2D, SeparableConv2D, DepthwiseConv2D, Dense, MaxPooling2D, AveragePooling2D, GlobalAveragePooling2D,
        Flatten, Dropout, BatchNormalization, Activation, ZeroPadding2D, Add, Multiply, Concatenate)})
    layers.Conv1D, layers.SeparableConv1D, layers.Input = Conv1D, SeparableConv1D, Input
    models = mod("keras.models", Model=Model, Sequential=object)
    optimizers = mod("keras.optimizers", SGD=object, Adam=object)

    def no_multi_gpu(*a, **k):
        raise Exception("single device")

    utils = mod("keras.utils", multi_gpu_model=no_multi_gpu)
    mod("keras", backend=backend, layers=layers, models=models, optimizers=optimizers, utils=utils)
    device_lib = mod("tensorflow.python.client.device_lib", list_local_devices=lambda: [])
    client = mod("tensorflow.python.client", device_lib=device_lib)
    python = mod("tensorflow.python", client=client)
    mod("tensorflow", python=python)
    mod("validators")


# ----------------------------------------------------------------------------- cases
SHAPES = {"mnist": (28, 28, 1), "cifar": (32, 32, 3)}


def record(model, shape):
    net = model.build(shape, 10)
    if not net:
        return {"valid": False}
    return {"valid": True, "params": net.count_params(), "layers": net.describe()}


def main():
    if not os.path.isdir(REFERENCE):
        sys.exit("reference tree not found at %s" % REFERENCE)
    _install_stubs()
    sys.path.insert(0, REFERENCE)
    import numpy as np
    from model.keras_model import KerasFeatureModel
    from model.mutation.mutable_base import MutableBase, MutationStrategies
    from Products_tree import ProductSet

    golden = {"templates": {}, "products": {}, "mutation_walks": {}, "breeding": {}, "reparse": {}}
    quiet = open(os.devnull, "w")
    real_stdout = sys.stdout
    sys.stdout = quiet
    try:
        # A. templates
        for tpl in ("lenet5", "keras"):
            for ds, shape in SHAPES.items():
                model = KerasFeatureModel.parse_feature_model(tpl)
                golden["templates"]["%s@%s" % (tpl, ds)] = record(model, shape)

        # B. product files (the shipped one + the committed sampled fixtures)
        pdts = [os.path.join(REFERENCE, "datasets", "10Products.pdt")]
        pdts += sorted(os.path.join(HERE, f) for f in os.listdir(HERE) if f.endswith(".pdt"))
        for path in pdts:
            key = os.path.basename(path)
            rows = []
            pset = ProductSet(path)
            for product, original in pset.format_products():
                model = KerasFeatureModel.parse_feature_model(product, product_features=original)
                rows.append(record(model, SHAPES["cifar"]))
            golden["products"][key] = rows

        # C. seeded mutation walks (CHOICE strategy, the one every driver sets)
        MutableBase.mutation_stategy = MutationStrategies.CHOICE
        for tpl in ("lenet5", "keras"):
            for seed in range(6):
                np.random.seed(seed)
                model = KerasFeatureModel.parse_feature_model(tpl)
                steps = []
                for _ in range(12):
                    model.mutate(0.1)
                    steps.append(record(model, SHAPES["cifar"]))
                golden["mutation_walks"]["%s/seed%d" % (tpl, seed)] = steps

        # D. breeding (deep copies of block prefixes) followed by one mutation
        for seed in range(3):
            np.random.seed(100 + seed)
            a = KerasFeatureModel.parse_feature_model("lenet5")
            b = KerasFeatureModel.parse_feature_model("keras")
            for _ in range(3):
                a.mutate(0.1)
                b.mutate(0.1)
            rows = []
            for ratio in (1.0, 0.5, 0.0):
                child = a.breed(b, ratio)
                rows.append(record(child, SHAPES["cifar"]))
                child.mutate(0.1)
                rows.append(record(child, SHAPES["cifar"]))
            golden["breeding"]["seed%d" % seed] = rows

        # E. the in-place child reversal: depth=2 and parsing the same product dict twice
        path = os.path.join(HERE, "nas_5_5_10_seed0.pdt")
        if os.path.isfile(path):
            pset = ProductSet(path)
            rows = []
            for product, original in pset.format_products()[:4]:
                first = KerasFeatureModel.parse_feature_model(product)
                again = KerasFeatureModel.parse_feature_model(product)
                deep = KerasFeatureModel.parse_feature_model(copy.deepcopy(product), depth=2)
                rows.append([record(first, SHAPES["cifar"]), record(again, SHAPES["cifar"]),
                             record(deep, SHAPES["cifar"])])
            golden["reparse"]["nas_5_5_10_seed0.pdt"] = rows
    finally:
        sys.stdout = real_stdout
        quiet.close()

    out = os.path.join(HERE, "structural.json")
    with open(out, "w") as handle:
        json.dump(golden, handle, sort_keys=True, separators=(",", ":"))
    n_valid = sum(1 for rows in golden["products"].values() for r in rows if r["valid"])
    n_all = sum(len(rows) for rows in golden["products"].values())
    print("wrote %s (%d bytes); products valid %d/%d" % (out, os.path.getsize(out), n_valid, n_all))


if __name__ == "__main__":
    main()
