"""Ensembles of models (reference: veles/ensemble/)."""
