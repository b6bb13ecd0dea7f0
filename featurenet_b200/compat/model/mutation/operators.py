# -*- coding: utf-8 -*-
"""Mutation and breeding operators of the architecture description (pure host code).

Behavioural mirror of /root/reference/model/mutation/*.py.  Every operator draws from
``numpy.random`` with the same calls in the same order as the reference (``choice`` with /
without ``p``, ``rand``), so a run seeded with ``numpy.random.seed(s)`` walks through the
same mutants.  The as-shipped oddities are kept because the evolution drivers depend on the
states they create (last bullet):

* ``mutate_remove_block`` does nothing once a model has a block (mutable_model.py:39-41);
* a mutated input copies ``_kernel/_type/_activation`` from its predecessor even when the
  predecessor never had them (mutable_input.py:34-36);
* kernels up to 7x7 are written un-clamped (mutable_input.py:10,52-55);
* new cells / blocks are ``Cell.base_cell()`` / ``Block.base_block()``.
"""
import copy
from enum import Enum
from math import floor, ceil

from numpy.random import choice, rand


class MutationStrategies(Enum):
    CHOICE = 1   # one child per node is picked and mutated
    ALL = 2      # every child mutates with probability `rate`


class SelectionStrategies(Enum):
    ELITIST = 1
    HYBRID = 2
    PARETO = 3


class MutableBase(object):
    mutation_operators = []
    debug_mode = False
    mutation_stategy = MutationStrategies.CHOICE     # (sic) name used by run.py / full_evolution.py
    MAX_NB_CELLS = 10
    MAX_NB_BLOCKS = 20
    selection_stragey = SelectionStrategies.HYBRID    # (sic)

    @staticmethod
    def _pick_one():
        return MutableBase.mutation_stategy == MutationStrategies.CHOICE

    def mutate(self, rate=1):
        """mutable_base.py:22-36: weighted pick of one operator, or all of them in turn."""
        names, weights = zip(*self.mutation_operators)
        todo = [choice(names, None, weights)] if MutableBase._pick_one() else names
        for op_name in todo:
            outcome = getattr(self, op_name)(rate)
            if outcome and MutableBase.debug_mode:
                print("mutation {}".format(outcome))

    def _mutate_listed_attributes(self, rate, tag, draw):
        """Shared body of the ``mutate_attributes`` operators: re-draw one (or each) attribute."""
        done = []
        if MutableBase._pick_one():
            which = choice(list(self.attributes.keys()), None)
            value = draw(which)
            setattr(self, self.attributes[which], value)
            done = [(tag, which, value)]
        else:
            for which in self.attributes.keys():
                if rand() < rate:
                    value = draw(which)
                    setattr(self, self.attributes[which], value)
                    done.append((tag, which, value))
        return done

    def _swap_for_new(self, rate, candidates, install):
        """Shared body of the ``mutate_type`` operators: replace self by a fresh sibling."""
        if rand() < rate or MutableBase._pick_one():
            fresh = choice(candidates, None)()
            fresh.parent_cell = self.parent_cell
            return install(fresh)
        return None


# --------------------------------------------------------------------------- model level
class MutableModel(MutableBase):
    def __init__(self):
        self.mutation_operators = [("mutate_add_block", 0.3), ("mutate_block", 0.5), ("mutate_remove_block", 0.2)]

    def mutate_add_block(self, rate=1, block=None):
        from ..block import Block
        if len(self.blocks) == MutableBase.MAX_NB_BLOCKS:
            return ("add_block", None)
        draw = rand()
        if MutableBase._pick_one() or draw < rate:
            block = Block.base_block()
            self.blocks.insert(choice(range(len(self.blocks))), block)
            return ("add_block", block)

    def mutate_block(self, rate=1, block_index=None):
        if len(self.blocks) == 0:
            return ("mutate_block", None)
        if block_index is not None:
            return self.blocks[block_index].mutate()
        if MutableBase._pick_one():
            return self.mutate_block(rate, choice(len(self.blocks)))
        for idx in range(len(self.blocks)):
            self.mutate_block(rate, idx)

    def mutate_remove_block(self, rate=1, block_index=None):
        if len(self.blocks) >= 1:          # as shipped: never removes (mutable_model.py:39-41)
            return ("remove_block", None)
        if block_index is None:
            if MutableBase._pick_one():
                return self.mutate_remove_block(rate, choice(len(self.blocks)))
            for idx in range(len(self.blocks)):
                if rand() < rate:
                    self.mutate_remove_block(rate, idx)
        elif 0 <= block_index < len(self.blocks):
            del self.blocks[block_index]
            return ("remove_block", block_index)

    def breed(self, parent2, ratio=1.0):
        """Child = leading blocks of self + leading blocks of parent2 (mutable_model.py:56-63)."""
        from ..keras_model import KerasFeatureModel
        mine = copy.deepcopy(self.dump_blocks()["blocks"])
        theirs = copy.deepcopy(parent2.dump_blocks()["blocks"])
        merged = mine[:floor(len(mine) * ratio)] + theirs[:ceil(len(theirs) * (1 - ratio))]
        return KerasFeatureModel.parse_blocks({"blocks": merged})


# --------------------------------------------------------------------------- block level
class MutableBlock(MutableBase):
    cells = []
    attributes = {"strides_values": None, "features_multiplier_values": None}
    strides_values = ("1x1", "2x2")
    features_multiplier_values = (800, 400, 200, 100, 50, 25)

    def __init__(self, *_, **__):
        self.mutation_operators = (("mutate_add_cell", 0.3), ("mutate_cell", 0.3),
                                   ("mutate_remove_cell", 0.1), ("mutate_block", 0.3))

    def mutate_add_cell(self, rate=1):
        from ..cell import Cell
        if len(self.cells) == MutableBase.MAX_NB_CELLS:
            return ("mutate_add_cell", None)
        if rand() < rate or MutableBase._pick_one():
            cell = Cell.base_cell()
            where = choice(range(len(self.cells))) if len(self.cells) else 0
            self.cells.insert(where, cell)
        return ("mutate_add_cell", cell)      # unbound when nothing was added, as in the reference

    def mutate_block(self, rate=1):
        def draw(which):
            value = choice(getattr(self, which), None)
            self.attributes[which](value)
            return value

        done = []
        if MutableBase._pick_one():
            which = choice(list(self.attributes.keys()), None)
            done.append(("mutate_block", which, draw(which)))
        else:
            for which in list(self.attributes.keys()):
                if rand() < rate:
                    done.append(("mutate_block", which, draw(which)))
        return done

    def mutate_cell(self, rate=1, cell_index=None):
        if len(self.cells) == 0:
            return ("mutate_cell", None)
        if cell_index is not None:
            self.cells[cell_index].mutate(rate)
        elif MutableBase._pick_one():
            return self.mutate_cell(rate, choice(len(self.cells)))
        else:
            for idx in range(len(self.cells)):
                self.mutate_cell(rate, idx)

    def mutate_remove_cell(self, rate=1, cell_index=None):
        if len(self.cells) == 0:
            return ("mutate_remove_cell", None)
        if cell_index is not None and 0 <= cell_index < len(self.cells):
            del self.cells[cell_index]
            return ("mutate_remove_cell", cell_index)
        if MutableBase._pick_one():
            return self.mutate_remove_cell(rate, choice(len(self.cells)))
        for idx in range(len(self.cells)):
            if rand() < rate:
                return self.mutate_remove_cell(rate, idx)


# --------------------------------------------------------------------------- cell level
class MutableCell(MutableBase):
    def __init__(self, *_, **__):
        self.mutation_operators = (("mutate_input1", 0.4), ("mutate_input2", 0.3), ("mutate_output", 0.3))

    def mutate_input1(self, rate=1):
        self.input1.mutate(rate)
        return ("cell", "input1")

    def mutate_input2(self, rate=1):
        self.input2.mutate(rate)
        return ("cell", "input2")

    def mutate_operation1(self, rate=1):
        self.operation1.mutate()

    def mutate_operation2(self, rate=1):
        self.operation2.mutate()

    def mutate_combination(self, rate=1):
        self.combination.mutate()

    def mutate_output(self, rate=1):
        self.output.mutate(rate)
        return ("cell", "output")


# --------------------------------------------------------------------------- leaves
class MutableInput(MutableBase):
    attributes = {"kernel_values": "_kernel", "pool_type_values": "_type",
                  "conv_type_values": "_type", "activation_values": "_activation"}
    strides_values = ((1, 1), (2, 2))
    kernel_values = ((1, 1), (3, 1), (1, 3), (3, 3), (5, 1), (1, 5), (5, 5), (7, 1), (1, 7), (7, 7))
    pool_type_values = ("max", "average", "global")
    conv_type_values = ("normal", "separable", "depthwise")
    activation_values = ("relu", "sigmoid", None)
    features_values = (None, 8, 16, 32, 64, 128, 256, 512, 1024, 2048)

    def __init__(self, *_, **__):
        self.mutation_operators = (("mutate_type", 0.5), ("mutate_attributes", 0.5))

    def mutate_type(self, rate=1):
        from ..input import ZerosInput, IdentityInput, ConvolutionInput
        kinds = [IdentityInput, ConvolutionInput]
        if self.parent_cell.input2 == self:      # only the second input may become Zeros
            kinds.append(ZerosInput)

        def install(fresh):
            for attr in self.attributes.values():
                setattr(fresh, attr, getattr(self, attr, None))
            if self.parent_cell.input1 == self:
                self.parent_cell.input1 = fresh
            else:
                self.parent_cell.input2 = fresh
            return ("mutate_input_type", fresh)

        return self._swap_for_new(rate, kinds, install) or ("mutate_input_type",)

    def mutate_attributes(self, rate=1):
        def draw(which):
            values = getattr(self, which)
            return values[choice(len(values))]
        return self._mutate_listed_attributes(rate, "mutate_input_attribute", draw)


class MutableOperation(MutableBase):
    attributes = {"dropout_values": "_value", "activation_method": "_method"}
    dropout_values = (0.7, 0.3, 0.5)
    activation_method = ("relu", "tanh", "sigmoid")

    def __init__(self, *_, **__):
        self.mutation_operators = (("mutate_type", 0.5), ("mutate_attributes", 0.5))

    def mutate_type(self, rate=1):
        from ..operation import Active, BatchNorm, Void, Drop

        def install(fresh):
            for attr in self.attributes.values():
                setattr(fresh, attr, getattr(self, attr))
            if self.parent_cell.operation1 == self:
                self.parent_cell.operation1 = fresh
            else:
                self.parent_cell.operation2 = fresh
            return ("mutate_operation_type", fresh)

        return self._swap_for_new(rate, (Active, BatchNorm, Void, Drop), install) or ("mutate_operation_type",)

    def mutate_attributes(self, rate=1):
        return self._mutate_listed_attributes(rate, "mutate_operation_attribute",
                                              lambda which: choice(getattr(self, which), None))


class MutableCombination(MutableBase):
    def __init__(self, *_, **__):
        self.mutation_operators = (("mutate_type", 1))     # (sic) flat tuple, mutable_combination.py:10

    def mutate_type(self, rate=1):
        from ..operation import Concat, Sum

        def install(fresh):
            for attr in self.attributes.values():
                setattr(fresh, attr, getattr(self, attr))
            self.parent_cell.combination = fresh
            return ("mutate_combination_type", fresh)

        return self._swap_for_new(rate, (Concat, Sum), install) or ("mutate_combination_type",)


class MutableOutput(MutableBase):
    attributes = {"cell_index_values": "_relativeCellIndex"}
    cell_index_values = (1, 2, 3)

    def __init__(self, *_, **__):
        self.mutation_operators = (("mutate_type", 0.5), ("mutate_attributes", 0.5))

    def mutate_type(self, rate=1):
        from ..output import OutBlock, OutCell

        def install(fresh):
            for attr in self.attributes.values():
                setattr(fresh, attr, getattr(self, attr))
            self.parent_cell.output = fresh
            return ("mutate_output_type", fresh)

        return self._swap_for_new(rate, (OutBlock, OutCell), install) or ("mutate_output_type",)

    def mutate_attributes(self, rate=1):
        return self._mutate_listed_attributes(rate, "mutate_output_attribute",
                                              lambda which: choice(getattr(self, which), None))
