# -*- coding: utf-8 -*-
"""Base class of every element of a FeatureNet architecture description.

Mirrors /root/reference/model/node.py:3-78: a node carries an id-prefixed name, the raw
product sub-tree it was parsed from and the list of parameters the product left open.

One deliberate difference: the reference derives ``uniqid`` from ``uuid.uuid1()``, whose
first ten characters are the low 32 bits of a 100 ns clock, so "sort by name"
(keras_model.py:203, block.py:112) means "sort by creation time" except when that counter
wraps (every 429 s).  Here ``uniqid`` is a process-wide creation counter in fixed-width
hex: the same order without the wrap.
"""
import itertools

_creation_counter = itertools.count(1)


def type_of(element, keep_index=True):
    """Feature type = text after the last underscore of the label, lower-cased (node.py:61-69)."""
    label = element.get("label")
    kind = label[label.rfind("_") + 1:]
    if not keep_index:
        kind = "".join(ch for ch in kind if not ch.isdigit())
    return kind.lower()


class Node(object):
    content = None
    parent = None
    parent_model = None
    layer_mapping = {}
    node_mapping = {}

    def __init__(self, raw_dict=None, parent_model=None):
        self.parent_name = ""
        self.raw_dict = raw_dict
        self.customizable_parameters = {}
        self.uniqid = "%010x" % next(_creation_counter)
        Node.node_mapping[self.name] = self
        if parent_model:
            self.parent_model = parent_model

    # ---- naming
    def get_name(self, raw_dict=None):
        if self.raw_dict:
            label = self.raw_dict.get("label")
            for old, new in (("Block", "B"), ("Cell_", ""), ("Element", "C")):
                label = label.replace(old, new)
            return self.uniqid + label
        return self.uniqid

    @property
    def name(self):
        return self.get_name()

    # ---- open parameters
    def append_parameter(self, attribute, possible_values=""):
        self.customizable_parameters[attribute] = possible_values

    def get_custom_parameters(self):
        if self.customizable_parameters:
            return {self.get_name(): (self, self.customizable_parameters)}
        return {}

    # ---- layer <-> node bookkeeping
    @staticmethod
    def layer_to_node(layer):
        return Node.node_mapping.get(Node.layer_mapping.get(layer.name))

    @staticmethod
    def layer_to_cell(layer):
        return Node.layer_to_node(layer).parent_cell

    @staticmethod
    def node_to_layer(node):
        return [lay for (lay, n) in Node.layer_mapping.items() if n == node.name][0]

    get_type = staticmethod(type_of)

    # ---- ownership
    @property
    def parent_cell(self):
        return self._parent_cell

    @parent_cell.setter
    def parent_cell(self, value):
        self._parent_cell = value
        if value:
            self.parent_model = value.parent_model


def unwrap(x):
    """A stack entry is either a tensor or an Output holding one (input.py:31, operation.py:19)."""
    if hasattr(x, "content") and x.content is not None:
        return x.content
    return x


def selected_values(element):
    """{child type: type of its first selected grand-child} for an element of a product tree."""
    found = {}
    for child in element.get("children"):
        picks = child.get("children")
        if len(picks):
            found[type_of(child)] = type_of(picks[0])
    return found


def split_pair(text):
    """'3x1' -> ('3', '1'); anything that is not two fields -> None (input.py:94-99)."""
    if text is None:
        return None
    parts = text.split("x")
    return tuple(parts) if len(parts) == 2 else None
