# -*- coding: utf-8 -*-
"""A cell: two branches (input -> operation), a combination and an output routing.

Behaviour of /root/reference/model/cell.py:7-138, including the tensor-stack protocol of
``build_tensorflow_model`` (cell.py:47-86): a cell reads the entries of the block's stack
that are plain tensors or outputs whose countdown reached 0, and pushes its own result at
the head of the stack when its output is an ``OutCell``.
"""
from .mutation.mutable_cell import MutableCell
from .node import Node, type_of
from .input import Input, ZerosInput, IdentityInput, ConvolutionInput
from .output import Output, OutCell, OutBlock, Out  # noqa: F401
from .operation import Operation, Combination, Sum, Void

_PARSERS = {"input1": Input, "input2": Input, "operation1": Operation, "operation2": Operation,
            "combination": Combination, "output": Output}


class Cell(MutableCell, Node):
    def __init__(self, raw_dict=None, input1=None, operation1=None, input2=None, operation2=None,
                 output=None, output_combination=None):
        self._parent_block = None
        self.input1 = input1 if input1 else IdentityInput()
        self.input2 = input2 if input2 else ZerosInput()
        self.operation1 = operation1 if operation1 else Void()
        self.operation2 = operation2 if operation2 else Void()
        self.combination = output_combination if output_combination else Sum()
        self.output = output if output else OutCell()
        for part in (self.input1, self.input2, self.operation1, self.operation2, self.output, self.combination):
            part.parent_cell = self
        MutableCell.__init__(self)
        Node.__init__(self, raw_dict)

    @property
    def parent_block(self):
        return self._parent_block

    @parent_block.setter
    def parent_block(self, value):
        self._parent_block = value
        self.parent_model = value.parent_model

    def build_tensorflow_model(self, inputs, max_relative_index, block_stride, block_features):
        visible = [entry for entry in inputs if not hasattr(entry, "content") or entry.currentIndex == 0]

        if block_stride:
            self.input1.set_stride(block_stride)
            self.input2.set_stride(block_stride)
        if block_features:
            self.input1.set_features(block_features, True)
            self.input2.set_features(block_features, True)

        first = visible[0] if len(visible) > 0 else None
        i1 = self.input1.build(first)
        o1 = self.operation1.build(i1)
        Node.layer_mapping[i1.name] = self.input1.name
        Node.layer_mapping[o1.name] = self.operation1.name

        if type(self.input2) is ZerosInput:
            merged = o1                       # single-branch cell (cell.py:64-65)
        else:
            second = visible[1] if len(visible) > 1 else first
            i2 = self.input2.build(second, i1)
            i2.cell = self
            Node.layer_mapping[i2.name] = self.input2.name
            o2 = self.operation2.build(i2)
            Node.layer_mapping[o2.name] = self.operation2.name
            merged = self.combination.build(o1, o2)
        Node.layer_mapping[merged.name] = self.combination.name

        self.output.currentIndex = min(max_relative_index, self.output.currentIndex)
        routed = self.output.build(merged)
        Node.layer_mapping[routed.name] = self.output.name

        network_outputs = []
        if type(self.output) is OutCell and self.output.currentIndex > -1:
            inputs.insert(0, routed)
        if type(self.output) is Out:
            network_outputs.insert(0, routed.content)
        return network_outputs

    def get_custom_parameters(self):
        params = dict(Node.get_custom_parameters(self))
        for part in (self.input1, self.input2, self.operation1, self.operation2, self.combination):
            params.update(part.get_custom_parameters())
        return params

    @staticmethod
    def base_cell():
        """What mutations insert: one 3x3 convolution with 8 features and relu (cell.py:98-102)."""
        return Cell(input1=ConvolutionInput((3, 3), (1, 1), 8, "same", "relu"))

    @staticmethod
    def parse_feature_model(feature_model):
        feature_model["children"] = list(reversed(feature_model["children"]))   # in place, as the reference
        cell = Cell(raw_dict=feature_model)
        for part_dict in feature_model.get("children"):
            role = type_of(part_dict)
            part = None
            if len(part_dict.get("children")) > 0 and role in _PARSERS:
                part = _PARSERS[role].parse_feature_model(part_dict)
            if part:
                setattr(cell, role, part)
                setattr(part, "parent_cell", cell)
        return cell
