"""Reproducible PRNGs (reference: veles/prng/)."""
from .random_generator import RandomGenerator, get, xorshift128plus  # noqa: F401
