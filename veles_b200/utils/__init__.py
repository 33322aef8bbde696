"""Utilities: clocks sampling, timing helpers."""
