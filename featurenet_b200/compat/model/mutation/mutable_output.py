"""Import location kept from the reference (model/mutation/mutable_output.py); code lives in operators.py."""
from .operators import MutableOutput  # noqa: F401
