"""Mutation operators of the architecture description (see operators.py)."""
