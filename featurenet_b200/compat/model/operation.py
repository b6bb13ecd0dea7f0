# -*- coding: utf-8 -*-
"""Second stage of a cell branch (Operation) and the merge of the two branches (Combination).

Behaviour of /root/reference/model/operation.py:9-248.  Two hard-wired axes matter for the
kernels underneath: BatchNormalization runs with axis=1 and Concatenate with axis=1, i.e.
over / along the *height* of the NHWC tensors (operation.py:142,228).
"""
from .node import Node, type_of, unwrap
from .output import Output, OutCell, OutBlock, Out  # noqa: F401
from .mutation.mutable_operation import MutableOperation
from .mutation.mutable_combination import MutableCombination

from featurenet_b200 import layers as L


class Operation(MutableOperation, Node):
    def __init__(self, raw_dict=None, cell=None):
        self.parent_cell = cell
        MutableOperation.__init__(self)
        Node.__init__(self)

    def build_tensorflow_model(self, model, source1, source2):
        pass

    def build(self, input, neighbour=None):
        tensor = unwrap(input)
        Node.layer_mapping[tensor.name] = self.name
        return tensor

    @staticmethod
    def parse_feature_model(feature_model):
        chosen = feature_model.get("children")[0]
        kind = type_of(chosen)
        if kind == "void":
            return Void(chosen)
        if kind == "flatten":
            return Flat(chosen)
        if kind == "dropout":
            value = None
            for child in chosen.get("children"):
                if type_of(child) == "value" and len(child.get("children")):
                    value = type_of(child.get("children")[0])
            return Drop(_value=value, raw_dict=chosen)
        if kind == "padding":
            # "fillValue"/"fillSize" can never equal a lower-cased type: both stay unset (operation.py:46-58)
            return Padding(_fillValue=None, _fillSize=None, raw_dict=chosen)
        if kind == "batchnormalization":
            axis = None
            for child in chosen.get("children"):
                if type_of(child) == "axis" and len(child.get("children")):
                    axis = type_of(child.get("children")[0])
            return BatchNorm(_axis=axis, raw_dict=chosen)
        if kind == "activation":
            method = None
            for child in chosen.get("children"):
                if type_of(child) == "activation" and len(child.get("children")):
                    method = type_of(child.get("children")[0])
            return Active(method, raw_dict=chosen)
        return None


class Flat(Operation):
    def __init__(self, raw_dict=None, cell=None):
        super(Flat, self).__init__(raw_dict=raw_dict, cell=cell)

    def build(self, input):
        tensor = super(Flat, self).build(input)
        return L.Flatten()(tensor) if tensor.shape.ndims > 2 else tensor


class Void(Operation):
    def __init__(self, raw_dict=None, cell=None):
        super(Void, self).__init__(raw_dict=raw_dict, cell=cell)

    def build(self, input):
        return input


class Drop(Operation):
    def __init__(self, _value, raw_dict=None, cell=None):
        super(Drop, self).__init__(raw_dict=raw_dict, cell=cell)
        # product values are percentages: "5" -> 0.05; Drop(0.25) -> int(0.25)/100 == 0 (operation.py:107-110)
        self._value = int(_value) / 100 if _value else 0

    def build(self, input):
        tensor = super(Drop, self).build(input)
        return L.Dropout(self._value)(tensor) if self._value else tensor


class Padding(Operation):
    def __init__(self, _fillValue=None, _fillSize=None, raw_dict=None, cell=None):
        super(Padding, self).__init__(raw_dict=raw_dict, cell=cell)
        self._fillValue = 0
        self._fillSize = (int(_fillSize[0]), int(_fillSize[1])) if _fillSize else (1, 1)

    def build(self, input):
        tensor = super(Padding, self).build(input)
        return L.ZeroPadding2D(self._fillSize)(tensor) if tensor.shape.ndims == 4 else tensor


class BatchNorm(Operation):
    def __init__(self, _axis=None, raw_dict=None, cell=None):
        super(BatchNorm, self).__init__(raw_dict=raw_dict, cell=cell)
        self._axis = 1      # whatever the product says (operation.py:142)

    def build(self, input):
        tensor = super(BatchNorm, self).build(input)
        return L.BatchNormalization(axis=self._axis)(tensor)


class Active(Operation):
    _METHODS = ("tanh", "relu", "sigmoid", "softmax")

    def __init__(self, _method=None, raw_dict=None, cell=None):
        super(Active, self).__init__(raw_dict=raw_dict, cell=cell)
        _method = _method if _method else "relu"
        if str(_method) not in self._METHODS:
            self.append_parameter("_method", "|".join(self._METHODS))
        else:
            self._method = str(_method)

    def build(self, input):
        tensor = super(Active, self).build(input)
        return L.Activation(activation=self._method)(tensor)


class Combination(MutableCombination, Node):
    def __init__(self, raw_dict=None, cell=None):
        self.parent_cell = cell
        MutableCombination.__init__(self)
        Node.__init__(self)

    def build(self, source1, source2):
        """Align channel counts with a 1x1 "Reg_" convolution on the wider tensor (operation.py:173-188)."""
        self.parent_name = "{}+{}".format(source1.name, source2.name) if hasattr(source1, "name") else ""
        source1, source2 = unwrap(source1), unwrap(source2)
        if source1.shape.ndims == 4 and source2.shape.ndims == 4 \
                and source1.shape.dims[3].value != source2.shape.dims[3].value:
            if not source1.shape.dims[3] < source2.shape.dims[3]:
                source1, source2 = source2, source1
            # source1 now has fewer channels; source2 is regularised towards it
            dilation = 1
            stride = source1.shape.dims[2].value / source2.shape.dims[2].value
            if stride < 1:
                dilation = 1 / stride
                stride = 1
            source2 = L.Conv2D(source1.shape.dims[3].value, 1, strides=int(stride), dilation_rate=int(dilation),
                               name="Reg_" + Node.get_name(self))(source2)
        return source1, source2

    @staticmethod
    def parse_feature_model(feature_model):
        chosen = feature_model.get("children")[0]
        kind = type_of(chosen)
        if kind == "sum":
            return Sum(chosen)
        if kind == "product":
            return Product(chosen)
        if kind == "concat":
            axis = None
            for child in chosen.get("children"):
                if type_of(child) == "axis" and len(child.get("children")):
                    axis = type_of(child.get("children")[0])
            return Concat(_axis=axis, raw_dict=chosen)
        return None


def _same_shape(a, b):
    return a.shape.as_list() == b.shape.as_list()


class Sum(Combination):
    def __init__(self, raw_dict=None, cell=None):
        super(Sum, self).__init__(raw_dict=raw_dict, cell=cell)

    def build(self, source1, source2):
        source1, source2 = super(Sum, self).build(source1, source2)
        if _same_shape(source1, source2):
            return L.Add()([source1, source2])
        return source2          # shapes still differ: the second operand alone survives (operation.py:223)


class Concat(Combination):
    def __init__(self, _axis=None, raw_dict=None, cell=None):
        super(Concat, self).__init__(raw_dict=raw_dict, cell=cell)
        self._axis = 1

    def build(self, source1, source2):
        source1, source2 = super(Concat, self).build(source1, source2)
        if _same_shape(source1, source2):
            return L.Concatenate(axis=self._axis)([source1, source2])
        return source2


class Product(Combination):
    def __init__(self, raw_dict=None, cell=None):
        super(Product, self).__init__(raw_dict=raw_dict, cell=cell)

    def build(self, source1, source2):
        source1, source2 = super(Product, self).build(source1, source2)
        return L.Multiply()([source1, source2])
