"""Parallel execution: flat CUDA-graph step scheduler, synchronous data parallelism with the
fused all-reduce + gradient-descent kernel over NVLink/NVSwitch, NCCL baseline."""
from .step import FlatStep  # noqa: F401
