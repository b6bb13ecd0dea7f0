"""Placeholder kept for import compatibility (the reference's model/mobileNet.py defines no blocks)."""
