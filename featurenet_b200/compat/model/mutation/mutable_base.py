"""Import location kept from the reference (model/mutation/mutable_base.py); code lives in operators.py."""
from .operators import MutableBase, MutationStrategies, SelectionStrategies  # noqa: F401
