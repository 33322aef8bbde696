"""Genetic hyper-parameter search (reference: veles/genetics/)."""
from .config import Range, Tuneable, fix_config, process_config  # noqa: F401
from .core import Chromosome, Population  # noqa: F401
