"""Import location kept from the reference (model/mutation/mutable_model.py); code lives in operators.py."""
from .operators import MutableModel  # noqa: F401
