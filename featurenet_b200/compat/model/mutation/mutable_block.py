"""Import location kept from the reference (model/mutation/mutable_block.py); code lives in operators.py."""
from .operators import MutableBlock  # noqa: F401
