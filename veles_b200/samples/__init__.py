"""Samples that ship with the core (reference: ``veles/samples``); the neural network samples live in
``veles_b200/znicz/samples``."""
