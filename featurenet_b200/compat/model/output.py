# -*- coding: utf-8 -*-
"""Where a cell's result goes: next cells (OutCell), next block (OutBlock) or the network (Out).

Behaviour of /root/reference/model/output.py:7-95.  An ``OutCell`` result is parked on the
block's tensor stack and becomes visible ``relativeCellIndex + 1`` cells later; the
countdown lives in ``currentIndex`` (decremented by Block after every cell, block.py:58-60).
"""
from .node import Node, type_of, unwrap
from .mutation.mutable_output import MutableOutput


class Output(MutableOutput, Node):
    currentIndex = 0
    _relativeBlockIndex = None
    _relativeCellIndex = None

    def __init__(self, raw_dict=None, cell=None):
        self.parent_cell = cell
        MutableOutput.__init__(self)
        Node.__init__(self)          # the reference's MRO drops raw_dict on this path

    def __deepcopy__(self, memo):
        # tensors parked in `content` must not be cloned with the description
        twin = type(self)()
        twin._relativeBlockIndex = self._relativeBlockIndex
        twin._relativeCellIndex = self._relativeCellIndex
        twin.parent_cell = self.parent_cell
        return twin

    @property
    def shape(self):
        return self.content.shape if self.content is not None else None

    def build(self, input):
        self.content = unwrap(input)
        return self

    def build_tensorflow_model(self, model, source1, source2):
        pass

    @staticmethod
    def parse_feature_model(feature_model):
        chosen = feature_model.get("children")[0]
        kind = type_of(chosen)
        if kind == "block":
            # get_type lower-cases, so "relativeBlockIndex" is never recognised (output.py:47)
            return OutBlock(_relativeBlockIndex=None, raw_dict=chosen)
        if kind == "cell":
            index = None
            for child in chosen.get("children"):
                if type_of(child) == "relativecellindex" and len(child.get("children")):
                    index = type_of(child.get("children")[0])
            return OutCell(_relativeCellIndex=index, raw_dict=chosen)
        return Out(chosen)


class Out(Output):
    def __init__(self, raw_dict=None, cell=None):
        super(Out, self).__init__(raw_dict=raw_dict, cell=cell)


class OutBlock(Output):
    def __init__(self, _relativeBlockIndex=None, raw_dict=None, cell=None):
        super(OutBlock, self).__init__(raw_dict=raw_dict, cell=cell)
        self._relativeBlockIndex = int(_relativeBlockIndex) if _relativeBlockIndex else 0
        self.currentIndex = self._relativeBlockIndex


class OutCell(Output):
    def __init__(self, _relativeCellIndex=None, raw_dict=None, cell=None):
        super(OutCell, self).__init__(raw_dict=raw_dict, cell=cell)
        # the product value is a string: "0" is truthy and becomes 1 (output.py:84-88)
        self._relativeCellIndex = int(_relativeCellIndex) + 1 if _relativeCellIndex else 1

    def build(self, input):
        super(OutCell, self).build(input)
        self.currentIndex = self._relativeCellIndex
        return self
