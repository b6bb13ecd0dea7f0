"""The "keras" template (block list of /root/reference/model/kerasNet.py:7-33).

Kept with its as-shipped wiring: both max-pool cells are appended to the *first* block and
``Drop(0.25)`` evaluates to no dropout (int(0.25)/100 == 0).
"""
from .input import PoolingInput, ConvolutionInput, DenseInput
from .operation import Drop
from .block import Block
from .cell import Cell


def _conv(features):
    return Cell(input1=ConvolutionInput((3, 3), (1, 1), features, "same", "relu"))


def _pool():
    return Cell(input1=PoolingInput((2, 2), (1, 1), "max", "valid"), operation1=Drop(0.25))


def standard_blocks():
    block1, block2, block4 = Block(), Block(), Block()
    block1.append_cell(_conv(32))
    block1.append_cell(_conv(32))
    block1.append_cell(_pool())
    block2.append_cell(_conv(64))
    block2.append_cell(_conv(64))
    block1.append_cell(_pool())          # (sic) kerasNet.py:25 appends to block1
    block4.append_cell(Cell(input1=DenseInput(512, "relu"), operation1=Drop(0.25)))
    return [block1, block2, block4]
