# -*- coding: utf-8 -*-
"""A block: an ordered list of cells sharing a stride and a feature multiplier.

Behaviour of /root/reference/model/block.py:7-113.  Worth knowing when reading
``build_tensorflow_model``: the reference filters the stack with ``i is OutBlock`` (an
instance compared with a class, always false), so a block's result is always rebuilt from
the head of the stack and cells routed to ``OutBlock`` never feed anything (block.py:67-73).
"""
from .mutation.mutable_block import MutableBlock
from .node import Node
from .cell import Cell
from .output import Out, OutCell, OutBlock  # noqa: F401


def _suffix(label):
    return label[label.rfind("_") + 1:]


class Block(MutableBlock, Node):
    def __init__(self, raw_dict=None, previous_block=None, parent_model=None):
        self.is_root = True
        self.cells = []
        self._features = 0
        self._stride = 0
        if previous_block:
            self.previous_block = previous_block
            self.is_root = False
        self.attributes = {"strides_values": self.set_stride, "features_multiplier_values": self.set_features}
        MutableBlock.__init__(self)
        Node.__init__(self, raw_dict, previous_block)     # (sic) block.py:25 -> mutable_base.py:40

    def append_cell(self, cell):
        cell.parent_block = self
        self.cells.append(cell)

    def set_stride(self, stride_str):
        self._stride = stride_str.split("x")

    def set_features(self, features_str):
        self._features = int(features_str) / 100

    def get_custom_parameters(self):
        params = dict(Node.get_custom_parameters(self))
        for cell in self.cells:
            params.update(cell.get_custom_parameters())
        return params

    def build_tensorflow_model(self, inputs):
        block_input = inputs[0]
        outputs = []
        nb_cells = len(self.cells)
        for position, cell in enumerate(self.cells):
            outputs = outputs + cell.build_tensorflow_model(inputs, nb_cells - position, self._stride, self._features)
            # parked cell outputs move one cell closer to being visible
            for entry in inputs:
                if type(entry) is OutCell and entry.currentIndex >= 0:
                    entry.currentIndex = entry.currentIndex - 1

        carried = [entry for entry in inputs if entry is OutBlock]       # (sic) always empty
        if len(carried) == 0:
            result = OutBlock()
            result.parent_name = self.get_name()
            result.build(inputs[0])
            carried = [result]
        carried.append(block_input)
        return carried, outputs

    @staticmethod
    def base_block():
        block = Block()
        block.append_cell(Cell.base_cell())
        return block

    @staticmethod
    def parse_feature_model(feature_model, parent_model=None):
        block = Block(raw_dict=feature_model, parent_model=parent_model)
        for child in feature_model.get("children"):
            role = _suffix(child.get("label"))
            if role == "stride":
                block.set_stride(_suffix(child.get("children")[0].get("label")))
            elif role == "features":
                block.set_features(_suffix(child.get("children")[0].get("label")))
            elif len(child.get("children")):
                inner = child.get("children")[0]
                inner_kind = "".join(ch for ch in _suffix(inner.get("label")) if not ch.isdigit())
                if inner_kind == "Cell":
                    block.append_cell(Cell.parse_feature_model(inner))
        block.cells.sort(key=lambda c: c.get_name())
        return block
