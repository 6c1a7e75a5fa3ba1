"""Modules named after the reference files they shadow (utils/*.py)."""
