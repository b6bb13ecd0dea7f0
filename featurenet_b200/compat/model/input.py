# -*- coding: utf-8 -*-
"""First stage of a cell branch: the operation applied to the tensor a cell reads.

Behaviour of /root/reference/model/input.py:11-308.  The clamps and coercions are the
reference's: conv kernels are clipped to 5 per side and pool kernels to 3 at parse time,
"valid" is not an accepted padding so everything runs "same", an unknown conv/pool type
becomes "normal"/"max", block strides are clamped to 1..2, and feature counts derived from a
block multiplier are clamped to 6..2048.
"""
from .node import Node, type_of, unwrap, selected_values, split_pair
from .output import Output, OutCell, OutBlock, Out  # noqa: F401  (import surface of the reference)
from .mutation.mutable_input import MutableInput

from featurenet_b200 import layers as L


def _rank(t):
    return t.shape.ndims


class Input(MutableInput, Node):
    min_features = 6
    max_features = 2048

    def __init__(self, raw_dict=None, stride=1, features=0, cell=None):
        self._stride = stride
        self._features = features
        self._relative_features = None
        self.parent_cell = cell
        MutableInput.__init__(self)
        Node.__init__(self)          # raw_dict is lost on this path in the reference too

    def set_stride(self, stride):
        sx, sy = int(stride[0]), int(stride[1])
        self._stride = (max(1, min(sx, 2)), max(1, min(sy, 2)))

    def set_features(self, features, relative_features=False):
        if relative_features:
            self._relative_features = float(features)
        else:
            self._features = max(Input.min_features, min(int(features), Input.max_features))

    def build(self, input, neighbour=None):
        tensor = unwrap(input)
        if self._relative_features:
            self.set_features(tensor.shape.as_list()[-1] * self._relative_features)
        if not self._features:
            self._features = Input.min_features
        return tensor

    @staticmethod
    def parse_feature_model(feature_model):
        chosen = feature_model.get("children")[0]
        kind = type_of(chosen)
        if kind == "zeros":
            return ZerosInput(raw_dict=chosen)
        if kind == "identity":
            return IdentityInput(raw_dict=chosen)
        if kind not in ("dense", "pooling", "convolution"):
            return None
        picked = selected_values(chosen)
        if kind == "dense":
            return DenseInput(_features=picked.get("features"), _activation=picked.get("activation"), raw_dict=chosen)
        geometry = dict(_kernel=split_pair(picked.get("kernel")), _stride=split_pair(picked.get("stride")),
                        _type=picked.get("type"), _padding=picked.get("padding"), raw_dict=chosen)
        if kind == "pooling":
            return PoolingInput(**geometry)
        return ConvolutionInput(_activation=picked.get("activation", "relu"), _features=picked.get("features"), **geometry)


class ZerosInput(Input):
    def __init__(self, raw_dict=None, cell=None):
        super(ZerosInput, self).__init__(raw_dict=raw_dict, cell=cell)

    def build(self, input, neighbour=None):
        # K.zeros on a shape with an unknown batch axis cannot be materialised (input.py:163-165);
        # only reached when Zeros is the *first* input, which constraint C7 forbids.
        raise ValueError("Cannot convert a partially known TensorShape to a Tensor")


class IdentityInput(Input):
    def __init__(self, raw_dict=None, cell=None):
        super(IdentityInput, self).__init__(raw_dict=raw_dict, cell=cell)

    def build(self, input, neighbour=None):
        return input


_ACTIVATIONS = ("tanh", "relu", "sigmoid", "softmax", "none")


class DenseInput(Input):
    def __init__(self, _features=128, _activation="relu", raw_dict=None, cell=None):
        super(DenseInput, self).__init__(raw_dict=raw_dict, cell=cell)
        if not _features:
            self.append_parameter("_features", "__int__")
        else:
            self._features = int(_features)
        if not _activation or str(_activation) not in _ACTIVATIONS:
            self.append_parameter("_activation", "|".join(_ACTIVATIONS))
        else:
            self._activation = str(_activation) if _activation != "none" else None

    def build(self, input, neighbour=None):
        tensor = super(DenseInput, self).build(input)
        return L.Dense(self._features, activation=self._activation, name=Node.get_name(self))(tensor)


class PoolingInput(Input):
    _TYPES = ("max", "average", "global")
    _PADDINGS = ("same",)

    def __init__(self, _kernel=(3, 3), _stride=(1, 1), _type="max", _padding="same", raw_dict=None, cell=None):
        super(PoolingInput, self).__init__(raw_dict=raw_dict, cell=cell)
        if not _type or str(_type) not in self._TYPES:
            self.append_parameter("_type", "|".join(self._TYPES))
            self._type = "max"
        else:
            self._type = _type
        if self._type == "global":
            return
        # a missing value leaves the attribute unset: _kernel then fails at build time (input.py:208-211)
        if not _kernel:
            self.append_parameter("_kernel", "(__int__,__int__)")
        else:
            self._kernel = (min(int(_kernel[0]), 3), min(int(_kernel[1]), 3))
        if not _stride:
            self.append_parameter("_stride", "(__int__,__int__)")
        else:
            self._stride = (int(_stride[0]), int(_stride[1]))
        if not _padding or str(_padding) not in self._PADDINGS:
            self.append_parameter("_padding", "|".join(self._PADDINGS))
            self._padding = self._PADDINGS[0]
        else:
            self._padding = _padding

    def _fix_valid(self, tensor):
        dims = tensor.shape.dims
        if self._padding == "valid" and (self._kernel[0] > dims[1].value or self._kernel[1] > dims[2].value):
            self._padding = "same"

    def build(self, input, neighbour=None):
        tensor = unwrap(super(PoolingInput, self).build(input))
        if _rank(tensor) != 4:
            return tensor
        if self._type in ("max", "average"):
            self._fix_valid(tensor)
            layer = L.MaxPooling2D if self._type == "max" else L.AveragePooling2D
            return layer(pool_size=self._kernel, strides=self._stride, padding=self._padding,
                         name=Node.get_name(self))(tensor)
        if self._type == "global":
            return L.GlobalAveragePooling2D(name=Node.get_name(self))(tensor)
        return tensor


class ConvolutionInput(Input):
    _TYPES = ("normal", "separable", "depthwise")
    _PADDINGS = ("same",)

    def __init__(self, _kernel=(3, 3), _stride=(1, 1), _features=8, _padding="same", _activation="relu",
                 _type="normal", raw_dict=None, cell=None):
        super(ConvolutionInput, self).__init__(raw_dict=raw_dict, cell=cell)
        if not _type or str(_type) not in self._TYPES:
            self.append_parameter("_type", "|".join(self._TYPES))
            self._type = "normal"
        else:
            self._type = _type
        if not _kernel:
            self.append_parameter("_kernel", "(__int__,__int__)")
        else:
            self._kernel = (min(int(_kernel[0]), 5), min(int(_kernel[1]), 5))
        if not _stride:
            self.append_parameter("_stride", "(__int__,__int__)")
        else:
            self._stride = (int(_stride[0]), int(_stride[1]))
        if not _features:
            self.append_parameter("_features", "__int__")
        else:
            self._features = int(_features)
        if not _activation or str(_activation) not in _ACTIVATIONS:
            self.append_parameter("_activation", "|".join(_ACTIVATIONS))
        else:
            self._activation = str(_activation) if _activation != "none" else None
        if not _padding or str(_padding) not in self._PADDINGS:
            self.append_parameter("_padding", "|".join(self._PADDINGS))
            self._padding = self._PADDINGS[0]
        else:
            self._padding = _padding

    def build(self, input, neighbour=None):
        tensor = unwrap(super(ConvolutionInput, self).build(input))
        rank = _rank(tensor)

        def common():
            return dict(strides=self._stride, padding=self._padding, activation=self._activation,
                        name=Node.get_name(self))

        if rank == 4:
            if self._type == "normal":
                dims = tensor.shape.dims
                if self._padding == "valid" and (self._kernel[0] > dims[1].value or self._kernel[1] > dims[2].value):
                    self._padding = "same"
                return L.Conv2D(self._features, self._kernel, **common())(tensor)
            if self._type == "separable":
                return L.SeparableConv2D(self._features, self._kernel, **common())(tensor)
            if self._type == "depthwise":
                return L.DepthwiseConv2D(self._kernel, **common())(tensor)
        elif rank == 3:
            if self._type == "normal":
                return L.Conv1D(self._features, self._kernel[0], **common())(tensor)
            if self._type == "separable":
                return L.SeparableConv1D(self._features, self._kernel[0], **common())(tensor)
        # any other type (e.g. "max" written by a mutation) leaves the tensor untouched (input.py:291-308)
        return tensor
