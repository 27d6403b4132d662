"""Placeholder (see package docstring)."""
