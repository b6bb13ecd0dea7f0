"""Import location kept from the reference (model/mutation/mutable_operation.py); code lives in operators.py."""
from .operators import MutableOperation  # noqa: F401
