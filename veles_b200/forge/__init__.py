"""Forge: the model zoo (reference: veles/forge/)."""
