"""Import location kept from the reference (model/mutation/mutable_cell.py); code lives in operators.py."""
from .operators import MutableCell  # noqa: F401
