"""Import location kept from the reference (model/mutation/mutable_input.py); code lives in operators.py."""
from .operators import MutableInput  # noqa: F401
