"""Znicz — the neural network engine (re-specified: the reference's veles/znicz submodule is
absent from the checkout)."""
