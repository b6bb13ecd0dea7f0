"""Import location kept from the reference (model/mutation/mutable_combination.py); code lives in operators.py."""
from .operators import MutableCombination  # noqa: F401
