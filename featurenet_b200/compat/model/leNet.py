"""The "lenet5" template (block list of /root/reference/model/leNet.py:7-41).

As emitted: 5x5 tanh convs of 6*Cin, 12 and 120 features, 2x2 average pools of stride 1 and
2 (every "valid" is coerced to "same"), Dense 84 tanh applied to the 4-D tensor.
"""
from .input import PoolingInput, ConvolutionInput, DenseInput
from .block import Block
from .cell import Cell


def _block(stride, *first_inputs, **kw):
    block = Block()
    block.set_stride(stride)
    if "features" in kw:
        block.set_features(kw["features"])
    for first in first_inputs:
        block.append_cell(Cell(input1=first))
    return block


def lenet5_blocks():
    return [
        _block("1x1", ConvolutionInput((5, 5), None, None, "same", "tanh"),
               PoolingInput((2, 2), None, "average", "valid"), features=600),
        _block("1x1", ConvolutionInput((5, 5), None, 12, "same", "tanh")),
        _block("2x2", PoolingInput((2, 2), None, "average", "valid")),
        _block("1x1", ConvolutionInput((5, 5), (1, 1), 120, "valid", "tanh")),
        _block("1x1", DenseInput(84, "tanh")),
    ]
